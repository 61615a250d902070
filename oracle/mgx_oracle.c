/* TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the MGARD-X hot path.
 * See mgx_oracle.h.  Build: gcc -std=c11 -O2 -ffp-contract=off -shared -fPIC.
 * Parity pin: checked against oracle/_ref (the reference compiled from
 * /root/reference) in tests/test_oracle_vs_ref.py and against tests/golden/.
 */
#define _GNU_SOURCE
#include "mgx_oracle.h"

#include <float.h>
#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* Hierarchy::init level shapes, reference Hierarchy/Hierarchy.hpp:277-309:
 * per dim n -> n/2+1 while n > 2, then 2; l_target = min_d(#levels_d) - 1;
 * level_shape[l][d] = shape_level[d][l_target - l]. */
int mgxo_levels(int D, const uint64_t *shape, uint64_t max_level,
                uint64_t *level_shapes) {
  uint64_t sl[MGXO_MAX_D][64];
  int cnt[MGXO_MAX_D];
  int nlevel = INT_MAX;
  for (int d = 0; d < D; d++) {
    uint64_t n = shape[d];
    int k = 0;
    while (n > 2) {
      sl[d][k++] = n;
      n = n / 2 + 1;
    }
    sl[d][k++] = 2;
    cnt[d] = k;
    if (k < nlevel)
      nlevel = k;
  }
  (void)cnt;
  uint64_t L = (uint64_t)(nlevel - 1);
  if (max_level < L)
    L = max_level;
  for (uint64_t l = 0; l <= L; l++)
    for (int d = 0; d < D; d++)
      level_shapes[l * D + d] = sl[d][L - l];
  return (int)L;
}

/* Grid of the hierarchies built from now on (mgxo_set_grid): the caller's node
 * coordinates -- Hierarchy(shape, coords, config), Hierarchy.hpp:840-857, data
 * structure Cartesian_Grid_Non_Uniform -- or a uniform grid on the unit cube /
 * with unit spacing (Config::normalize_coordinates).  Test infrastructure: process
 * wide, not thread safe. */
static int mgxo_grid_nonuniform = 0;
static int mgxo_grid_normalize = 1;
static const void *mgxo_grid_coords[MGXO_MAX_D] = {0};

void mgxo_set_grid(int D, const void *const *coords, int normalize_coordinates) {
  mgxo_grid_nonuniform = coords != NULL;
  mgxo_grid_normalize = normalize_coordinates != 0;
  for (int d = 0; d < MGXO_MAX_D; d++)
    mgxo_grid_coords[d] = (coords && d < D) ? coords[d] : NULL;
}

#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)

#define T float
#define SUF(x) CAT(x, _f32)
#define SQRT_T sqrtf
#define FABS_T fabsf
#define COPYSIGN_T copysignf
#define EXP2_T exp2f
#define EPS_T FLT_EPSILON
#include "mgx_oracle_core.inc"
#undef T
#undef SUF
#undef SQRT_T
#undef FABS_T
#undef COPYSIGN_T
#undef EXP2_T
#undef EPS_T

#define T double
#define SUF(x) CAT(x, _f64)
#define SQRT_T sqrt
#define FABS_T fabs
#define COPYSIGN_T copysign
#define EXP2_T exp2
#define EPS_T DBL_EPSILON
#include "mgx_oracle_core.inc"
#undef T
#undef SUF

#include "mgx_oracle_huffman.inc"
