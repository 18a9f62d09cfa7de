// cumsum_dev.cu -- placeholder until the tile-parallel exact scan lands.
#include "context.hpp"

namespace amclb200
{
void exact_cumsum_parallel(PfCtx *c, const Particle *set, int n, double *d_cum)
{
  (void)c; (void)set; (void)n; (void)d_cum;
  AMCL_FATAL("exact_cumsum_parallel not built");
}
}  // namespace amclb200
