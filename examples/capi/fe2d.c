/* An application written against the reference's C interface (include/cedar/capi.h: bmg2_topo_create,
 * bmg2_operator_create/set/apply, bmg2_solver_create/run) -- plain C, no C++ and no CUDA in sight.  It enters the
 * 9-point finite-element Laplacian (centre 8, eight neighbours -1) VERTEX by VERTEX with signed entries, as an
 * application would (the interface converts to BoxMG's symmetric storage, src/2d/interface/c/operator.cc:30-58),
 * checks the operator application against the stencil, solves A x = b with b = A * x_exact, and prints
 * "max |A x - stencil|", "max |x - x_exact|".  Usage: fe2d [n]   (run where ./config.json configures the solver). */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <cedar/capi.h>

int main(int argc, char **argv)
{
	unsigned int n = argc > 1 ? (unsigned int)atoi(argv[1]) : 65;
	unsigned int lnx[1] = {n}, lny[1] = {n};
	bmg2_topo topo = bmg2_topo_create(0, n, n, lnx, lny, 1, 1);
	if (!topo) return 2;
	bmg2_operator op = bmg2_operator_create(topo);
	if (!op) return 2;

	/* all 9 entries of every vertex, neighbours outside the grid left out */
	static const int di[9] = {0, -1, 0, -1, -1, 1, 0, 1, 1}, dj[9] = {0, 0, -1, -1, 1, -1, 1, 1, 0};
	static const bmg2_dir dirs[9] = {BMG2_C, BMG2_W, BMG2_S, BMG2_SW, BMG2_NW, BMG2_SE, BMG2_N, BMG2_NE, BMG2_E};
	size_t cap = (size_t)n * n * 9, nv = 0;
	grid_coord_2d *coords = malloc(cap * sizeof(*coords));
	double *vals = malloc(cap * sizeof(*vals));
	for (unsigned int j = 0; j < n; j++)
		for (unsigned int i = 0; i < n; i++)
			for (int d = 0; d < 9; d++) {
				int ii = (int)i + di[d], jj = (int)j + dj[d];
				if (ii < 0 || jj < 0 || ii >= (int)n || jj >= (int)n) continue;
				coords[nv].i = i; coords[nv].j = j; coords[nv].dir = dirs[d];
				vals[nv] = d == 0 ? 8.0 : -1.0;
				nv++;
			}
	bmg2_operator_set(op, (unsigned int)nv, coords, vals);

	size_t N = (size_t)n * n;
	double *xe = malloc(N * sizeof(double)), *b = malloc(N * sizeof(double)), *x = calloc(N, sizeof(double));
	for (unsigned int j = 0; j < n; j++)
		for (unsigned int i = 0; i < n; i++) xe[(size_t)j * n + i] = sin(0.37 * (i + 1)) * cos(0.11 * (j + 1)) + 0.01 * i;
	bmg2_operator_apply(op, xe, b);
	double aerr = 0;
	for (unsigned int j = 0; j < n; j++)
		for (unsigned int i = 0; i < n; i++) {
			double s = 8.0 * xe[(size_t)j * n + i];
			for (int d = 1; d < 9; d++) {
				int ii = (int)i + di[d], jj = (int)j + dj[d];
				if (ii < 0 || jj < 0 || ii >= (int)n || jj >= (int)n) continue;
				s -= xe[(size_t)jj * n + ii];
			}
			aerr = fmax(aerr, fabs(s - b[(size_t)j * n + i]));
		}

	bmg2_solver slv = bmg2_solver_create(&op);
	if (!slv) return 3;
	bmg2_solver_run(slv, x, b);
	double xerr = 0;
	for (size_t k = 0; k < N; k++) xerr = fmax(xerr, fabs(x[k] - xe[k]));
	printf("n %u apply_err %.3e solve_err %.3e\n", n, aerr, xerr);
	bmg_timer_save("timer.json");
	bmg2_solver_destroy(slv);
	bmg2_operator_destroy(op);
	bmg2_topo_destroy(topo);
	free(coords); free(vals); free(xe); free(b); free(x);
	return (aerr < 1e-12 && xerr < 1e-6) ? 0 : 1;
}
