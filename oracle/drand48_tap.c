/*
 * drand48_tap.c -- TEST INFRASTRUCTURE (oracle side only; never linked into
 * the product library).
 *
 * A drand48()/srand48() pair that is bit-identical to glibc's (48-bit LCG,
 * X' = 0x5DEECE66D * X + 0xB mod 2^48, srand48(s) => X = (s << 16) | 0x330E,
 * result = X / 2^48) and that can log every uniform it hands out.  It is
 * linked INTO oracle/_ref/libamcl_ref.so with -Bsymbolic so that the
 * unmodified reference objects (pf.c:55,430,460; pf_pdf.c:68,139-141) draw
 * from it, which lets a test (a) seed the reference deterministically,
 * (b) count draws, (c) replay the exact stream elsewhere.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static uint64_t tap_state = 0x1234ABCD330EULL;
static uint64_t tap_draws = 0;
static int tap_ignore_srand = 0;

static double *tap_log = 0;
static uint64_t tap_log_cap = 0;
static uint64_t tap_log_len = 0;

double drand48(void)
{
  tap_state = (0x5DEECE66DULL * tap_state + 0xBULL) & 0xFFFFFFFFFFFFULL;
  double u = (double)tap_state * (1.0 / 281474976710656.0); /* exact: 48 bits */
  tap_draws++;
  if (tap_log && tap_log_len < tap_log_cap)
    tap_log[tap_log_len++] = u;
  return u;
}

void srand48(long int seed)
{
  if (tap_ignore_srand)
    return;
  tap_state = ((((uint64_t)seed) & 0xFFFFFFFFULL) << 16) | 0x330EULL;
}

/* ---- controls for the test harness ---- */

void tap_seed(long int seed)
{
  tap_state = ((((uint64_t)seed) & 0xFFFFFFFFULL) << 16) | 0x330EULL;
}

/* When set, the reference's own srand48(time(NULL)) / srand48(++seed) calls
 * are swallowed and only tap_seed() moves the stream. */
void tap_set_ignore_srand(int flag) { tap_ignore_srand = flag; }

uint64_t tap_get_state(void) { return tap_state; }
void tap_set_state(uint64_t s) { tap_state = s & 0xFFFFFFFFFFFFULL; }

uint64_t tap_draw_count(void) { return tap_draws; }
void tap_reset_count(void) { tap_draws = 0; }

void tap_log_begin(uint64_t capacity)
{
  free(tap_log);
  tap_log = (double *)malloc(sizeof(double) * (capacity ? capacity : 1));
  tap_log_cap = capacity;
  tap_log_len = 0;
}

uint64_t tap_log_size(void) { return tap_log_len; }

void tap_log_copy(double *out, uint64_t n)
{
  if (n > tap_log_len)
    n = tap_log_len;
  memcpy(out, tap_log, sizeof(double) * n);
}

void tap_log_end(void)
{
  free(tap_log);
  tap_log = 0;
  tap_log_cap = tap_log_len = 0;
}
