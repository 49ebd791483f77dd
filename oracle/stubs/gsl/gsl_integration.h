// oracle/stubs/gsl/gsl_integration.h -- STAND-IN for GSL's integration header (test infrastructure only), so that the
// reference's cgfunc/FixedPointQuadrature.h and cgfunc/Quadrature.h compile unmodified.  GSL (>= 2.0, un-vendored,
// version unpinned: README.md:39) is absent from this image; what is restated here from its published sources is
//   gsl_integration_glfixed_table_alloc / gsl_integration_glfixed  (integration/glfixed.c): m = (n+1)/2 non-negative
//     abscissae ascending (x0 = 0 for odd n), A = (b-a)/2, B = (b+a)/2,
//     odd n:  s = w0 f(B) + sum_{i>=1} w_i (f(B + A x_i) + f(B - A x_i)),  even n: the same sum from i = 0;  return A s
//   gsl_integration_qag(key = 1)  (integration/qag.c, qk15.c, ...): handed to the oracle's own restatement, orc_gsl_qag15.
// Nodes and weights come from oracle/gl_tables.h (40-digit recomputation, oracle/gen_gl_tables.py).  PARITY UNPINNED for
// everything in this file.
#pragma once
#include <cstddef>
#include <cstdlib>

struct gsl_function { double (*function)(double x, void* params); void* params; };

struct gsl_integration_glfixed_table { size_t n; double* x; double* w; int precomputed; };

extern "C" {
// implemented in oracle/ref_loops_shim.cpp on top of oracle/gl_tables.h
gsl_integration_glfixed_table* gsl_integration_glfixed_table_alloc(size_t n);
void gsl_integration_glfixed_table_free(gsl_integration_glfixed_table* t);
double gsl_integration_glfixed(const gsl_function* f, double a, double b, const gsl_integration_glfixed_table* t);

struct gsl_integration_workspace { size_t limit; };
gsl_integration_workspace* gsl_integration_workspace_alloc(size_t n);
void gsl_integration_workspace_free(gsl_integration_workspace* w);
int gsl_integration_qag(const gsl_function* f, double a, double b, double epsabs, double epsrel, size_t limit, int key,
                        gsl_integration_workspace* workspace, double* result, double* abserr);
}
