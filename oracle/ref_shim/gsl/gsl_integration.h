// TEST INFRASTRUCTURE: declarations only, so that the reference's Integrator.H (included by
// QMOM.C:41, whose code never instantiates integrate()) parses.  GSL is absent from this image.
#ifndef REF_SHIM_GSL_INTEGRATION_H
#define REF_SHIM_GSL_INTEGRATION_H
#include <stddef.h>
extern "C" {
typedef struct gsl_integration_workspace gsl_integration_workspace;
typedef struct { double (*function)(double x, void *params); void *params; } gsl_function;
gsl_integration_workspace *gsl_integration_workspace_alloc(size_t n);
void gsl_integration_workspace_free(gsl_integration_workspace *w);
int gsl_integration_qagiu(gsl_function *f, double a, double epsabs, double epsrel, size_t limit,
                          gsl_integration_workspace *workspace, double *result, double *abserr);
}
#endif
