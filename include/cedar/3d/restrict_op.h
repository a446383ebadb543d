// API surface after Cedar (c) 2016-2017 Cedar Developers, BSD-3-Clause -- see NOTICE at the repository root.
#ifndef CEDAR_3D_INTER_RESTRICT_OP_H
#define CEDAR_3D_INTER_RESTRICT_OP_H
#include <cedar/3d/prolong_op.h>
namespace cedar { namespace cdr3 {
// R = P^T: a handle on the prolongator (reference 3d/restrict_op.h)
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
