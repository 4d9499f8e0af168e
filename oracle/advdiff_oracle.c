/* TEST INFRASTRUCTURE ONLY -- see advdiff_oracle.h.  Build: `make -C oracle oracle`
 * (gcc -O3 -march=native -ffp-contract=off -fopenmp).  OpenMP only parallelises
 * independent lines / cells; results do not depend on the thread count. */
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include <math.h>
#include <stddef.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "advdiff_oracle.h"

#define FN2(name, S) name##_##S
#define FN1(name, S) FN2(name, S)

#define REAL double
#define FN(name) FN1(name, f64)
#include "advdiff_oracle_impl.inc"
#undef REAL
#undef FN

#define REAL float
#define FN(name) FN1(name, f32)
#include "advdiff_oracle_impl.inc"
#undef REAL
#undef FN

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; a timed CPU baseline has to ask for the
 * host's cores explicitly.  n <= 0 leaves the setting alone. */
void orc_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
