"""kISS-2's word model in numpy (test infrastructure): which LMS suffixes PREFIX_DOUBLING with a bounded k leaves tied.

kiss2_core.hpp:20-100 (encode_S) cuts every LMS substring into words of 16 bases — a word starts at the LMS position and
every 16 bases after it while it is before the next LMS position, and reads 16 bases whatever follows, zero past the end —
and the doubling loop (:764-785) runs index_offset = 1, 2, 4, ... < k over WORD indices, so after it two LMS suffixes are
in one group iff their first W = 2^ceil(log2 k) words are equal (words that overlap the text end are groups of their own,
:130-133, :215-217).  Inside such a group the reference's order is whatever std::sort left (:261-285): it depends on the
thread count, so outputs are compared after sorting every such run by position ("canonical",
SURVEY.md 8(c)(iii))."""
import numpy as np


def word_classes(S, lms, k):
    """class id per LMS suffix lms[0..m-1) (ascending text positions, the sentinel n excluded): equal ids <=> the first
    W words are equal, W = number of words the doubling loop covers for this k"""
    S = np.asarray(S, np.uint8)
    n = S.size
    lms = np.asarray(lms, np.int64)
    nxt = np.concatenate([lms[1:], [n]])
    cnt = (nxt - lms + 15) // 16                      # words per LMS substring (kiss2_core.hpp:45-47)
    first = np.concatenate([[0], np.cumsum(cnt)[:-1]])  # word index of every LMS start
    nw = int(cnt.sum())
    owner = np.repeat(np.arange(lms.size), cnt)
    start = lms[owner] + 16 * (np.arange(nw) - first[owner])
    Sp = np.concatenate([S, np.zeros(16, np.uint8)]).astype(np.uint64)
    val = np.zeros(nw, np.uint64)
    for j in range(16):
        val = (val << np.uint64(2)) | Sp[start + j]
    tail = start > n - 16                               # :81-84
    r = val.copy()                                      # initial groups: equal word value; tail words alone
    r[tail] = (np.uint64(1) << np.uint64(33)) + np.arange(int(tail.sum()), dtype=np.uint64)
    _, r = np.unique(r, return_inverse=True)
    r = r.astype(np.int64) + 1
    off = 1
    while off < k and off <= (2**32 - 1) // 2:          # :764-766
        key2 = np.zeros(nw, np.int64)
        if off < nw:
            key2[:nw - off] = r[off:]                   # get_key: 0 past the end (:102-111)
        _, r = np.unique(r * (nw + 2) + key2, return_inverse=True)
        r = r.astype(np.int64) + 1
        if np.unique(r).size == nw:
            break
        off *= 2
    return r[first]


def canonical(sorted_lms, S, k):
    """(`sorted_lms` (sentinel first) with every maximal run of suffixes of one word class put in ascending position,
    number of runs with more than one suffix)"""
    sorted_lms = np.asarray(sorted_lms, np.int64)
    n = np.asarray(S).size
    assert int(sorted_lms[0]) == n
    body = sorted_lms[1:]
    if body.size == 0:
        return sorted_lms.astype(np.uint32), 0
    lms = np.sort(body)
    cls = word_classes(S, lms, k)
    c = cls[np.searchsorted(lms, body)]
    run = np.concatenate([[0], np.cumsum(c[1:] != c[:-1])])
    order = np.lexsort((body, run))
    return np.concatenate([[n], body[order]]).astype(np.uint32), int((np.bincount(run) > 1).sum())
