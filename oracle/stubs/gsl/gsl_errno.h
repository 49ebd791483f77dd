// oracle/stubs/gsl/gsl_errno.h -- stand-in (test infrastructure): the reference's quadrature wrappers include it for
// nothing but the declarations GSL's integration header needs.
#pragma once
#define GSL_SUCCESS 0
