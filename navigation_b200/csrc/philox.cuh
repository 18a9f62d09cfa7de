// philox.cuh -- Philox4x32-10 counter-based generator (Salmon et al., SC'11),
// written out by hand so the library has no cuRAND dependency.  One stream per
// (key, particle): the counter is (particle index, block counter), so any
// thread can produce its own uniforms with no shared state -- the device-side
// replacement for the reference's single global drand48() stream
// (/root/reference/amcl/src/amcl/pf/pf_pdf.c:139-141, pf.c:430,460).
#pragma once

#include <stdint.h>

namespace amclb200
{

struct PhiloxStream
{
  uint32_t key[2];
  uint32_t ctr[4];
  uint32_t out[4];
  int have;  // unread 32-bit words left in out[]

  __host__ __device__ PhiloxStream(uint64_t k, uint64_t index)
  {
    key[0] = (uint32_t)k;
    key[1] = (uint32_t)(k >> 32);
    ctr[0] = (uint32_t)index;
    ctr[1] = (uint32_t)(index >> 32);
    ctr[2] = 0;
    ctr[3] = 0;
    have = 0;
  }

  __host__ __device__ static inline void mulhilo(uint32_t a, uint32_t b, uint32_t *hi, uint32_t *lo)
  {
    const uint64_t p = (uint64_t)a * (uint64_t)b;
    *hi = (uint32_t)(p >> 32);
    *lo = (uint32_t)p;
  }

  __host__ __device__ inline void refill()
  {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
#pragma unroll
    for (int r = 0; r < 10; r++)
    {
      uint32_t hi0, lo0, hi1, lo1;
      mulhilo(0xD2511F53u, c0, &hi0, &lo0);
      mulhilo(0xCD9E8D57u, c2, &hi1, &lo1);
      const uint32_t n0 = hi1 ^ c1 ^ k0;
      const uint32_t n2 = hi0 ^ c3 ^ k1;
      c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
      k0 += 0x9E3779B9u;
      k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
    have = 4;
    if (++ctr[2] == 0) ++ctr[3];
  }

  // Uniform double in [0, 1) with 53 random bits.
  __host__ __device__ inline double uniform()
  {
    if (have < 2) refill();
    const uint32_t a = out[4 - have], b = out[5 - have];
    have -= 2;
    const uint64_t bits = (((uint64_t)a << 32) | b) >> 11;
    return (double)bits * (1.0 / 9007199254740992.0);
  }

  // Zero-mean Gaussian by the polar Box-Muller rejection loop of
  // pf_ran_gaussian (pf_pdf.c:133-147).
  __host__ __device__ inline double gaussian(double sigma)
  {
    double x1, x2, w, r;
    do
    {
      do { r = uniform(); } while (r == 0.0);
      x1 = 2.0 * r - 1.0;
      do { r = uniform(); } while (r == 0.0);
      x2 = 2.0 * r - 1.0;
      w = x1 * x1 + x2 * x2;
    } while (w > 1.0 || w == 0.0);
    return sigma * x2 * sqrt(-2.0 * log(w) / w);
  }

  // Three Gaussians, value for value what three gaussian() calls in a row
  // return: pairs are drawn from the stream in the same order and the k-th
  // ACCEPTED pair makes the k-th Gaussian.  One merged rejection loop instead of
  // three keeps a warp's lanes together (the slowest lane of three separate loops
  // costs ~10 rounds, the slowest lane of the merged loop ~7), and the
  // log / sqrt / divide run once, outside any divergent code.
  __host__ __device__ inline void gaussian3(double s0, double s1, double s2, double g[3])
  {
    double xa = 0, wa = 1, xb = 0, wb = 1, xc = 0, wc = 1;
    int k = 0;
    while (k < 3)
    {
      double r, x1, x2, w;
      do { r = uniform(); } while (r == 0.0);
      x1 = 2.0 * r - 1.0;
      do { r = uniform(); } while (r == 0.0);
      x2 = 2.0 * r - 1.0;
      w = x1 * x1 + x2 * x2;
      if (w > 1.0 || w == 0.0) continue;
      if (k == 0) { xa = x2; wa = w; }
      else if (k == 1) { xb = x2; wb = w; }
      else { xc = x2; wc = w; }
      k++;
    }
    g[0] = s0 * xa * sqrt(-2.0 * log(wa) / wa);
    g[1] = s1 * xb * sqrt(-2.0 * log(wb) / wb);
    g[2] = s2 * xc * sqrt(-2.0 * log(wc) / wc);
  }
};

}  // namespace amclb200
