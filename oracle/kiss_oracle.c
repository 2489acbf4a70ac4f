/*
 * oracle/kiss_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C, CPU restatement of jhhung/kISS's `suffix_sort` path
 * (kiss1_core: get_lms -> lms_suffix_direct_sort_dna -> put_lms_suffix -> induced_sort),
 * written from the reference's behaviour, one function per reference function, each
 * citing the file:line it follows (paths relative to /root/reference/include/biovoltron/algo/sort/).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
 * load this library; the product (kiss_b200/csrc) never links or calls it.
 *
 * Parity status: PINNED.  The reference has no golden vectors for this path (tests/kiss.cpp
 * holds only a randomized k-order property), so the restatement is pinned against outputs of the
 * reference itself: oracle/_ref/libkiss_ref.so (the unmodified reference headers compiled by
 * oracle/build_ref.sh) — live in tests/test_oracle.py whenever that .so is present — and
 * against tests/golden/*.json fixtures generated from it by tests/golden/make_golden.py.
 *
 * All arithmetic is integer: u8 codes 0..3, u32 indices, EMPTY = 0xFFFFFFFF (constant.hpp:19-20).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define KISS_EMPTY 0xFFFFFFFFu
#define KISS_STRIDE_DNA 125u /* constant.hpp:29  KISS1_SPLIT_SORT_STRIDE_DNA */
#define KISS_PREFIX 10u      /* constant.hpp:27  KISS1_SPLIT_SORT_PREFIX_SIZE */

/* ------------------------------------------------------------------------------------------
 * get_lms — kiss_common.hpp:483-541 (per-thread scan) + 543-579 (driver).
 *
 * Right-to-left scan with a 2-bit type shift register: bit0 = type of the position just
 * examined (1 = S, 0 = L), bit1 = type of the position to its right.  Position i is an LMS
 * position iff register == 0b10 (type(i) = S, type(i-1) = L)  (constant.hpp:12-16, :516-522).
 * The last suffix is L-type (:503-504, ptr == n).  Output: LMS positions ascending, then the
 * sentinel n; returns m including the sentinel (:560-578).  The result does not depend on the
 * reference's thread count, so one scan over [0, n) restates all threads at once.
 *
 * count[c]     = number of positions holding symbol c         (bucket[4*256 + c], :529-536)
 * lms_count[c] = number of LMS positions holding symbol c     (bucket[LMS_TYPE*256 + c], :519)
 * ------------------------------------------------------------------------------------------ */
uint32_t oracle_get_lms(const uint8_t* S, uint32_t n, uint32_t* lms_out, uint32_t count[4],
                        uint32_t lms_count[4]) {
  for (int c = 0; c < 4; c++) count[c] = lms_count[c] = 0;
  if (n == 0) {
    lms_out[0] = 0;
    return 1;
  }
  uint32_t m = 0;
  uint8_t c0 = S[n - 1];
  uint8_t type = 0; /* L_TYPE: the run reaching the end of the text is L (:504) */
  for (uint32_t i = n - 1; i >= 1; i--) {
    count[c0]++;
    uint8_t c1 = S[i - 1];
    type = (uint8_t)((type << 1) + (c1 == c0 ? (type & 1) : (c1 < c0)));
    if ((type & 3) == 2) { /* LMS_TYPE */
      lms_count[c0]++;
      lms_out[m++] = i; /* descending for now */
    }
    c0 = c1;
  }
  count[c0]++; /* position 0 is never LMS (:535-536) */
  /* reverse to ascending text order and append the sentinel (:567-578) */
  for (uint32_t a = 0, b = m; a + 1 < b; a++, b--) {
    uint32_t t = lms_out[a];
    lms_out[a] = lms_out[b - 1];
    lms_out[b - 1] = t;
  }
  lms_out[m++] = n;
  return m;
}

/* ------------------------------------------------------------------------------------------
 * The LMS order — comparator lambda of lms_suffix_direct_sort_dna, kiss1_core.hpp:94-135.
 * Returns non-zero iff suffix i sorts before suffix j.  On packed text the reference compares
 * 125-base blocks as 256-bit big-endian values (structs.hpp:122-169); on unpacked codes that is
 * a 125-byte memcmp.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const uint8_t* S;
  uint32_t n, k;
} lms_cmp_ctx;

static int lms_less(const lms_cmp_ctx* cx, uint32_t i, uint32_t j) {
  const uint8_t* S = cx->S;
  const uint64_t n = cx->n, k = cx->k;
  uint64_t sorted_len = 0, a = i, b = j;
  while (sorted_len <= k && a + KISS_STRIDE_DNA <= n && b + KISS_STRIDE_DNA <= n) { /* :96-98 */
    int r = memcmp(S + a, S + b, KISS_STRIDE_DNA);                                   /* :104-114 */
    if (r != 0) return r < 0;
    sorted_len += KISS_STRIDE_DNA; /* :117-119 */
    a += KISS_STRIDE_DNA;
    b += KISS_STRIDE_DNA;
  }
  while (sorted_len + 1 <= k && a < n && b < n) { /* :121-130 */
    if (S[a] != S[b]) return S[a] < S[b];
    sorted_len++;
    a++;
    b++;
  }
  if (sorted_len >= k) return a < b; /* :131-133  position tie-break */
  return a == n;                     /* :134      the shorter suffix first */
}

int oracle_lms_less(const uint8_t* S, uint32_t n, uint32_t k, uint32_t i, uint32_t j) {
  lms_cmp_ctx cx = {S, n, k};
  return lms_less(&cx, i, j);
}

static int lms_qsort_cmp(const void* pa, const void* pb, void* ctx) {
  uint32_t a = *(const uint32_t*)pa, b = *(const uint32_t*)pb;
  if (a == b) return 0;
  return lms_less((const lms_cmp_ctx*)ctx, a, b) ? -1 : 1;
}

/* top `KISS_PREFIX` bases of suffix idx as an integer, bases past the end read as 0
 * (structs.hpp:175-184 load_prefix_length_less_than_16; padding: structs.hpp:94-96). */
static uint32_t prefix10(const uint8_t* S, uint32_t n, uint32_t idx) {
  uint32_t v = 0;
  for (uint32_t t = 0; t < KISS_PREFIX; t++) {
    uint64_t p = (uint64_t)idx + t;
    v = (v << 2) | (p < n ? S[p] : 0u);
  }
  return v;
}

/* ------------------------------------------------------------------------------------------
 * lms_suffix_direct_sort_dna — kiss1_core.hpp:23-145.
 * Stable counting sort of the m LMS suffixes (sentinel included) into 4^10 buckets by their
 * first 10 bases (:41-83), then a comparison sort inside every bucket (:137-142).
 * lms[0..m) in, sorted order out (in place); sorted[0] == n afterwards.
 * ------------------------------------------------------------------------------------------ */
void oracle_lms_suffix_direct_sort_dna(const uint8_t* S, uint32_t n, uint32_t* lms, uint32_t m,
                                       uint32_t k) {
  const uint32_t NB = 1u << (2 * KISS_PREFIX);
  uint32_t* start = (uint32_t*)calloc((size_t)NB + 1, sizeof(uint32_t));
  uint32_t* buf = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(m ? m : 1));
  for (uint32_t i = 0; i < m; i++) start[prefix10(S, n, lms[i]) + 1]++;
  for (uint32_t b = 0; b < NB; b++) start[b + 1] += start[b];
  uint32_t* cur = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)NB);
  memcpy(cur, start, sizeof(uint32_t) * (size_t)NB);
  for (uint32_t i = 0; i < m; i++) buf[cur[prefix10(S, n, lms[i])]++] = lms[i];
  lms_cmp_ctx cx = {S, n, k};
#pragma omp parallel for schedule(dynamic, 64)
  for (uint32_t b = 0; b < NB; b++) {
    uint32_t len = start[b + 1] - start[b];
    if (len > 1) qsort_r(buf + start[b], len, sizeof(uint32_t), lms_qsort_cmp, &cx);
  }
  memcpy(lms, buf, sizeof(uint32_t) * (size_t)m); /* :144 */
  free(cur);
  free(buf);
  free(start);
}

/* ------------------------------------------------------------------------------------------
 * put_lms_suffix — kiss_common.hpp:445-481.
 * SA[0..m) holds the sorted LMS suffixes (SA[0] = n) on entry.  On exit the sorted LMS suffixes
 * of symbol c occupy the last lms_count[c] slots of bucket c, everything else in 1..n is EMPTY.
 * ------------------------------------------------------------------------------------------ */
void oracle_put_lms_suffix(uint32_t* SA, uint32_t n, uint32_t m, const uint32_t count[4],
                           const uint32_t lms_count[4]) {
  uint32_t end[4], acc = 1; /* inclusive scan with initial value 1 (:463) */
  for (int c = 0; c < 4; c++) {
    acc += count[c];
    end[c] = acc;
  }
  uint32_t* SA1 = SA + 1;
  uint32_t m1 = m - 1;        /* SA1.size() (:465) */
  uint64_t j = (uint64_t)n + 1; /* SA.size()  (:466) */
  for (int c = 3; c >= 0; c--) {
    uint32_t num = lms_count[c];
    if (num == 0) continue;
    uint32_t i = end[c];
    for (uint64_t t = i; t < j; t++) SA[t] = KISS_EMPTY; /* :473-474 */
    j = i - num;
    m1 -= num;
    memmove(&SA[j], &SA1[m1], sizeof(uint32_t) * (size_t)num); /* :476 */
  }
  for (uint64_t t = 1; t < j; t++) SA[t] = KISS_EMPTY; /* :480 */
}

/* ------------------------------------------------------------------------------------------
 * induced_sort — kiss_common.hpp:422-442, with the sequential block bodies
 * induced_L_block :226-246 (run over every non-EMPTY slot, as induced_L :372-403 does),
 * induced_clear :405-420 and induced_S_block :40-61 (via induced_S :192-224).
 * The reference's parallel block variants are re-schedulings of exactly these loops.
 * ------------------------------------------------------------------------------------------ */
void oracle_induced_L(const uint8_t* S, uint32_t n, uint32_t* SA, const uint32_t count[4]) {
  uint32_t cur[4], acc = 1; /* get_L_bucket: exclusive scan, initial value 1 (:27-38) */
  for (int c = 0; c < 4; c++) {
    cur[c] = acc;
    acc += count[c];
  }
  for (uint64_t i = 0; i <= n; i++) {
    uint32_t v = SA[i];
    if (v == KISS_EMPTY || v == 0) continue; /* :383-387, :233 */
    uint32_t u = v - 1;
    uint8_t cv = (v == n) ? 0 : S[v]; /* :239 */
    uint8_t cu = S[u];
    SA[cur[cu]++] = (cu < cv) ? KISS_EMPTY : u; /* :241 (the EMPTY write is the reference's) */
  }
}

void oracle_induced_clear(uint32_t n, uint32_t* SA, const uint32_t count[4],
                          const uint32_t lms_count[4]) {
  (void)n;
  uint32_t acc = 1; /* get_S_bucket: inclusive scan, initial value 1 (:14-25) */
  for (int c = 0; c < 4; c++) {
    acc += count[c];
    for (uint32_t t = acc - lms_count[c]; t < acc; t++) SA[t] = KISS_EMPTY; /* :412-418 */
  }
}

void oracle_induced_S(const uint8_t* S, uint32_t n, uint32_t* SA, const uint32_t count[4]) {
  uint32_t cur[4], acc = 1;
  for (int c = 0; c < 4; c++) {
    acc += count[c];
    cur[c] = acc;
  }
  for (uint64_t i = n; i > 0; i--) { /* :202 block_beg from n down to 1 */
    uint32_t v = SA[i];
    if (v == KISS_EMPTY || v == 0) continue; /* :203-206, :50 */
    uint32_t u = v - 1;
    uint8_t cu = S[u], cv = S[v];
    if (cu <= cv) SA[--cur[cu]] = u; /* :55-56 */
  }
}

void oracle_induced_sort(const uint8_t* S, uint32_t n, uint32_t* SA, const uint32_t count[4],
                         const uint32_t lms_count[4]) {
  oracle_induced_L(S, n, SA, count);
  oracle_induced_clear(n, SA, count, lms_count);
  oracle_induced_S(S, n, SA, count);
}

/* ------------------------------------------------------------------------------------------
 * kiss1_suffix_array_dna — kiss1_core.hpp:229-268.  SA has n+1 slots.
 * Optional snapshots (any may be NULL): lms_out[m], sorted_out[m], placed_out[n+1].
 * Returns m.
 * ------------------------------------------------------------------------------------------ */
uint32_t oracle_kiss1_suffix_array_dna(const uint8_t* S, uint32_t n, uint32_t k, uint32_t* SA,
                                       uint32_t* lms_out, uint32_t* sorted_out,
                                       uint32_t* placed_out, uint32_t count_out[4],
                                       uint32_t lms_count_out[4]) {
  if (n == 0) { /* :237-238 */
    SA[0] = 0;
    return 1;
  }
  uint32_t count[4], lms_count[4];
  /* no two LMS positions are adjacent, so m <= n/2 + 1 and the list fits the front of SA */
  uint32_t m = oracle_get_lms(S, n, SA, count, lms_count);
  if (lms_out) memcpy(lms_out, SA, sizeof(uint32_t) * (size_t)m);
  oracle_lms_suffix_direct_sort_dna(S, n, SA, m, k);
  if (sorted_out) memcpy(sorted_out, SA, sizeof(uint32_t) * (size_t)m);
  oracle_put_lms_suffix(SA, n, m, count, lms_count);
  if (placed_out) memcpy(placed_out, SA, sizeof(uint32_t) * ((size_t)n + 1));
  oracle_induced_sort(S, n, SA, count, lms_count);
  if (count_out) memcpy(count_out, count, sizeof(count));
  if (lms_count_out) memcpy(lms_count_out, lms_count, sizeof(lms_count));
  return m;
}

/* put_lms_suffix + induced_sort from a caller-supplied sorted LMS list (sorted[0] == n);
 * used to test LMS-order models and the CUDA sort independently of the induction. */
void oracle_induce_from_sorted_lms(const uint8_t* S, uint32_t n, const uint32_t* sorted, uint32_t m,
                                   uint32_t* SA) {
  uint32_t count[4], lms_count[4];
  uint32_t* tmp = (uint32_t*)malloc(sizeof(uint32_t) * ((size_t)n / 2 + 2));
  uint32_t m2 = oracle_get_lms(S, n, tmp, count, lms_count);
  free(tmp);
  (void)m2;
  memcpy(SA, sorted, sizeof(uint32_t) * (size_t)m);
  oracle_put_lms_suffix(SA, n, m, count, lms_count);
  oracle_induced_sort(S, n, SA, count, lms_count);
}

/* FNV-1a over the little-endian bytes of a u32 array: the checksum the golden fixtures carry. */
uint64_t oracle_fnv1a_u32(const uint32_t* a, uint64_t len) {
  uint64_t h = 0xcbf29ce484222325ull;
  const uint8_t* p = (const uint8_t*)a;
  for (uint64_t i = 0; i < len * 4; i++) {
    h ^= p[i];
    h *= 0x100000001b3ull;
  }
  return h;
}

int oracle_abi_version(void) { return 1; }
