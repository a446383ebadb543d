// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_2D_INTER_RESTRICT_OP_H
#define CEDAR_2D_INTER_RESTRICT_OP_H
#include <iostream>
#include <cedar/2d/prolong_op.h>
namespace cedar { namespace cdr2 {
// R = P^T: just a handle on the prolongator (reference 2d/restrict_op.h:13-28)
class restrict_op
{
public:
	restrict_op() : P(nullptr) {}
	restrict_op(prolong_op *P) : P(P) {}
	void associate(prolong_op *P) { this->P = P; }
	prolong_op &getP() { return *P; }
	const prolong_op &getP() const { return *P; }
private:
	prolong_op *P;
};
}}
#endif
