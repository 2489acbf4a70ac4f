/*
 * oracle/general_oracle.c — TEST INFRASTRUCTURE ONLY (tests/ and nothing else; see kiss_oracle.c).
 *
 * Plain-C restatement of the reference's GENERAL-ALPHABET path, kiss1_suffix_array
 * (include/biovoltron/algo/sort/kiss1_core.hpp:270-311): get_lms -> lms_suffix_direct_sort (:153-226) ->
 * put_lms_suffix -> induced_sort, the same templates as the DNA path instantiated over bytes with 256 buckets
 * (kiss_common.hpp: ThreadState::bucket is 5 x 256 wide for either alphabet).
 *
 * The LMS order: for n < 10^8 the reference sorts with `small_cmp` (:208-212), a lexicographical compare of the first
 * min(k, remaining) symbols, through an UNSTABLE std::sort (:214-219) — the order of LMS suffixes that agree on their
 * first k symbols is unspecified there.  This restatement breaks those ties by text position and says so; where no such
 * tie exists (and for k >= n, the true suffix array) the result is unique and is the reference's.  For n >= 10^8 the
 * reference uses the 32-symbol block comparator (:170-206), restated below as oracle_general_less_big.
 *
 * Parity status: PINNED on the unique cases — tests/test_general.py compares this file with
 * oracle/_ref/libkiss_ref.so (ref_kiss1_suffix_array, the unmodified reference) on tie-free and k >= n inputs, and
 * tests/golden/general_golden.json carries digests generated from the reference.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define KISS_EMPTY 0xFFFFFFFFu
#define SIGMA 256
#define KISS_STRIDE 32u              /* constant.hpp: KISS1_SPLIT_SORT_STRIDE */
#define KISS_SMALL_THRES 100000000u  /* kiss1_core.hpp:214 */

/* get_lms over bytes — kiss_common.hpp:483-579 (see oracle_get_lms in kiss_oracle.c for the scan) */
uint32_t oracle_general_get_lms(const uint8_t* S, uint32_t n, uint32_t* lms_out, uint32_t* count /*256*/,
                                uint32_t* lms_count /*256*/) {
  memset(count, 0, sizeof(uint32_t) * SIGMA);
  memset(lms_count, 0, sizeof(uint32_t) * SIGMA);
  if (n == 0) {
    lms_out[0] = 0;
    return 1;
  }
  uint32_t m = 0;
  uint8_t c0 = S[n - 1];
  uint8_t type = 0;
  for (uint32_t i = n - 1; i >= 1; i--) {
    count[c0]++;
    uint8_t c1 = S[i - 1];
    type = (uint8_t)((type << 1) + (c1 == c0 ? (type & 1) : (c1 < c0)));
    if ((type & 3) == 2) {
      lms_count[c0]++;
      lms_out[m++] = i;
    }
    c0 = c1;
  }
  count[c0]++;
  for (uint32_t a = 0, b = m; a + 1 < b; a++, b--) {
    uint32_t t = lms_out[a];
    lms_out[a] = lms_out[b - 1];
    lms_out[b - 1] = t;
  }
  lms_out[m++] = n;
  return m;
}

typedef struct {
  const uint8_t* S;
  uint32_t n, k;
} gen_ctx;

/* small_cmp — kiss1_core.hpp:208-212: -1 / 0 / +1 on the first min(k, remaining) symbols, the shorter range first
 * when one is a prefix of the other (std::ranges::lexicographical_compare) */
static int small_cmp3(const gen_ctx* cx, uint32_t i, uint32_t j) {
  uint64_t li = (uint64_t)i + cx->k < cx->n ? cx->k : cx->n - i;
  uint64_t lj = (uint64_t)j + cx->k < cx->n ? cx->k : cx->n - j;
  uint64_t l = li < lj ? li : lj;
  int c = memcmp(cx->S + i, cx->S + j, (size_t)l);
  if (c) return c < 0 ? -1 : 1;
  return li < lj ? -1 : li > lj ? 1 : 0;
}

/* the block comparator — kiss1_core.hpp:170-206: whole 32-symbol blocks while sorted_len <= k and both suffixes have
 * a whole block left, then single symbols while sorted_len + 1 <= k, then `sorted_len >= k ? i < j : i == n` */
int oracle_general_less_big(const uint8_t* S, uint32_t n, uint32_t k, uint32_t i0, uint32_t j0) {
  uint64_t i = i0, j = j0, sorted_len = 0;
  while (sorted_len <= k && i + KISS_STRIDE <= n && j + KISS_STRIDE <= n) {
    int c = memcmp(S + i, S + j, KISS_STRIDE);
    if (c) return c < 0;
    sorted_len += KISS_STRIDE;
    i += KISS_STRIDE;
    j += KISS_STRIDE;
  }
  while (sorted_len + 1 <= k && i < n && j < n) {
    if (S[i] != S[j]) return S[i] < S[j];
    sorted_len++;
    i++;
    j++;
  }
  if (sorted_len >= k) return i < j;
  return i == n;
}

static int gen_qsort_cmp(const void* pa, const void* pb, void* ctx) {
  const gen_ctx* cx = (const gen_ctx*)ctx;
  uint32_t a = *(const uint32_t*)pa, b = *(const uint32_t*)pb;
  if (cx->n < KISS_SMALL_THRES) {
    int c = small_cmp3(cx, a, b);
    if (c) return c;
    return a < b ? -1 : a > b; /* the reference leaves this case to an unstable sort: position chosen here */
  }
  if (oracle_general_less_big(cx->S, cx->n, cx->k, a, b)) return -1;
  if (oracle_general_less_big(cx->S, cx->n, cx->k, b, a)) return 1;
  return 0;
}

/* lms_suffix_direct_sort — kiss1_core.hpp:153-226.  lms[0..m): ascending LMS positions then the sentinel n; on exit
 * lms[0] = n and lms[1..m) sorted (the driver rotates the sentinel to the front: get_lms leaves it last and
 * put_lms_suffix reads SA[1..m), :300-301) */
void oracle_general_lms_sort(const uint8_t* S, uint32_t n, uint32_t* lms, uint32_t m, uint32_t k) {
  gen_ctx cx = {S, n, k};
  memmove(lms + 1, lms, sizeof(uint32_t) * (size_t)(m - 1));
  lms[0] = n;
  qsort_r(lms + 1, m - 1, sizeof(uint32_t), gen_qsort_cmp, &cx);
}

/* put_lms_suffix + induced_sort over 256 buckets — kiss_common.hpp:445-481, 422-442 (bodies as in kiss_oracle.c) */
static void general_put_and_induce(const uint8_t* S, uint32_t n, uint32_t* SA, uint32_t m, const uint32_t* count,
                                   const uint32_t* lms_count) {
  uint32_t end[SIGMA], cur[SIGMA], acc = 1;
  for (int c = 0; c < SIGMA; c++) {
    acc += count[c];
    end[c] = acc;
  }
  uint32_t* SA1 = SA + 1;
  uint32_t m1 = m - 1;
  uint64_t j = (uint64_t)n + 1;
  for (int c = SIGMA - 1; c >= 0; c--) {
    uint32_t num = lms_count[c];
    if (num == 0) continue;
    uint32_t i = end[c];
    for (uint64_t t = i; t < j; t++) SA[t] = KISS_EMPTY;
    j = i - num;
    m1 -= num;
    memmove(&SA[j], &SA1[m1], sizeof(uint32_t) * (size_t)num);
  }
  for (uint64_t t = 1; t < j; t++) SA[t] = KISS_EMPTY;
  /* L pass */
  acc = 1;
  for (int c = 0; c < SIGMA; c++) {
    cur[c] = acc;
    acc += count[c];
  }
  for (uint64_t i = 0; i <= n; i++) {
    uint32_t v = SA[i];
    if (v == KISS_EMPTY || v == 0) continue;
    uint32_t u = v - 1;
    int cv = (v == n) ? -1 : S[v]; /* the sentinel is below every symbol */
    int cu = S[u];
    SA[cur[cu]++] = (cu < cv) ? KISS_EMPTY : u;
  }
  /* clear the LMS slots */
  acc = 1;
  for (int c = 0; c < SIGMA; c++) {
    acc += count[c];
    for (uint32_t t = acc - lms_count[c]; t < acc; t++) SA[t] = KISS_EMPTY;
  }
  /* S pass */
  acc = 1;
  for (int c = 0; c < SIGMA; c++) {
    acc += count[c];
    cur[c] = acc;
  }
  for (uint64_t i = n; i > 0; i--) {
    uint32_t v = SA[i];
    if (v == KISS_EMPTY || v == 0) continue;
    uint32_t u = v - 1;
    uint8_t cu = S[u], cv = S[v];
    if (cu <= cv) SA[--cur[cu]] = u;
  }
}

/* kiss1_suffix_array — kiss1_core.hpp:270-311.  SA has n + 1 slots.  Returns m. */
uint32_t oracle_kiss1_suffix_array(const uint8_t* S, uint32_t n, uint32_t k, uint32_t* SA) {
  if (n == 0) { /* :278-279 */
    SA[0] = 0;
    return 1;
  }
  uint32_t count[SIGMA], lms_count[SIGMA];
  uint32_t m = oracle_general_get_lms(S, n, SA, count, lms_count);
  oracle_general_lms_sort(S, n, SA, m, k);
  general_put_and_induce(S, n, SA, m, count, lms_count);
  return m;
}
