/*
 * oracle/hamming_oracle.c — TEST INFRASTRUCTURE, not product code.
 *
 * CPU restatement of what the reference's Hamming matching computes.  The arithmetic lives in a
 * third-party dependency that is not vendored under /root/reference: opencv-python, pinned at
 * 4.6.0.66 in the reference's requirements.txt:2 (cv::BFMatcher -> cv::batchDistance ->
 * hal::normHamming).  Call sites in the reference:
 *     custom_bruteforce_matcher.py:160-172   BFMatcher(NORM_HAMMING).knnMatch(k=2) + Lowe ratio
 *     custom_bruteforce_matcher.py:193-199   BFMatcher(NORM_HAMMING, crossCheck=True).match
 *
 * The published algorithm restated here (OpenCV modules/core/src/batch_distance.cpp):
 *   - knn, K=2: for each query row scan the train rows in ascending index; a candidate enters the
 *     sorted K-list only if its distance is STRICTLY smaller than the current K-th entry and is
 *     inserted behind every entry with distance <= its own.  Result: the first two rows of a
 *     stable ascending sort of the distance row (lowest train index first among equals).
 *   - crossCheck: nearest neighbour (K=1, same tie rule) in both directions; query q keeps
 *     train t iff t = nn_train(q) and nn_query(t) = q; output ascending q; distance as integer.
 *
 * Parity pin: tests/test_oracle_hamming.py checks this file against cv2.BFMatcher of the
 * installed opencv-python (4.13.0 in this image; 4.6.0.66 is not installable offline) on
 * tie-heavy random inputs, and against the committed fixtures in tests/golden/ that were
 * generated from cv2.BFMatcher by tests/golden/make_golden.py.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this.  Build: make -C oracle  (gcc, no dependencies).
 */
#include <limits.h>
#include <stdint.h>
#include <string.h>

static inline int ham256(const uint8_t* a, const uint8_t* b) {
  uint64_t x[4], y[4];
  memcpy(x, a, 32);
  memcpy(y, b, 32);
  return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) +
         __builtin_popcountll(x[2] ^ y[2]) + __builtin_popcountll(x[3] ^ y[3]);
}

/* out: nq x 4 int32 {idx1, d1, idx2, d2}; missing neighbours are -1. */
void oracle_hamming_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* out) {
  for (int i = 0; i < nq; ++i) {
    int dist[2] = {INT_MAX, INT_MAX};
    int idx[2] = {-1, -1};
    const uint8_t* qi = q + (size_t)i * 32;
    for (int j = 0; j < nt; ++j) {
      const int d = ham256(qi, t + (size_t)j * 32);
      if (d < dist[1]) {
        int k = 0; /* K-2 .. 0 with K = 2 */
        for (; k >= 0 && dist[k] > d; --k) {
          idx[k + 1] = idx[k];
          dist[k + 1] = dist[k];
        }
        idx[k + 1] = j;
        dist[k + 1] = d;
      }
    }
    out[4 * i + 0] = idx[0];
    out[4 * i + 1] = idx[0] < 0 ? -1 : dist[0];
    out[4 * i + 2] = idx[1];
    out[4 * i + 3] = idx[1] < 0 ? -1 : dist[1];
  }
}

/* Nearest neighbour with lowest-index tie break (K = 1). */
static void nn1(const uint8_t* a, int na, const uint8_t* b, int nb, int32_t* idx, int32_t* dist) {
  for (int i = 0; i < na; ++i) {
    int bd = INT_MAX, bi = -1;
    const uint8_t* ai = a + (size_t)i * 32;
    for (int j = 0; j < nb; ++j) {
      const int d = ham256(ai, b + (size_t)j * 32);
      if (d < bd) {
        bd = d;
        bi = j;
      }
    }
    idx[i] = bi;
    dist[i] = bd;
  }
}

/* crossCheck=True.  work: caller scratch of 2*(nq+nt) int32.  Writes up to nq matches
 * (ascending query index) and returns their number. */
int oracle_hamming_crosscheck(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* work,
                              int32_t* mq, int32_t* mt, int32_t* md) {
  int32_t* sidx = work;
  int32_t* sdist = work + nq;
  int32_t* tidx = work + 2 * (size_t)nq;
  int32_t* tdist = tidx + nt;
  nn1(q, nq, t, nt, sidx, sdist); /* batchDistance(src1, src2) */
  nn1(t, nt, q, nq, tidx, tdist); /* batchDistance(src2, src1) */
  int n = 0;
  for (int i = 0; i < nq; ++i) {
    const int j = sidx[i];
    if (j >= 0 && tidx[j] == i) {
      mq[n] = i;
      mt[n] = j;
      md[n] = sdist[i];
      ++n;
    }
  }
  return n;
}

/* Lowe ratio on knn2 records: keeps row i iff it has two neighbours and den*d1 < num*d2, the
 * exact integer form of `m.distance < 0.75 * n.distance` (num/den = 3/4). */
int oracle_ratio_survivors(const int32_t* best2, int nq, int num, int den, int32_t* mq, int32_t* mt,
                           int32_t* md) {
  int n = 0;
  for (int i = 0; i < nq; ++i) {
    const int32_t* r = best2 + 4 * (size_t)i;
    if (r[2] >= 0 && (int64_t)den * r[1] < (int64_t)num * r[3]) {
      mq[n] = i;
      mt[n] = r[0];
      md[n] = r[1];
      ++n;
    }
  }
  return n;
}
