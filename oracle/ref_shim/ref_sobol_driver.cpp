// extern "C" drivers around the UNMODIFIED vendored Burkardt Sobol generator (include/sobol/sobol.cpp),
// compiled where it lies into oracle/_ref/libref_sobol.so.  TEST INFRASTRUCTURE.
#include "sobol.hpp"

extern "C" {
// point number *seed of the dim_num-dimensional sequence; *seed is advanced (include/sobol/sobol.cpp:239)
void ref_i4_sobol(int dim_num, int *seed, float *quasi) { i4_sobol(dim_num, seed, quasi); }
void ref_i8_sobol(int dim_num, long long *seed, double *quasi) { i8_sobol(dim_num, seed, quasi); }
}
