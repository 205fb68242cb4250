// Host-side pieces of the drop-in library (no device code): SURVEY.md section 8(f) rank 4.
//
// b2asr_edit_distance -- the Levenshtein distance the reference takes from the third-party `editdistance` package
// (C++ there too; absent from this image) in eval.py:17-18 and src/postprocess.py:135-146 (cal_cer).  Sequences are
// passed as int32 ids (characters' code points, or the index of each distinct word), unit costs, two-row DP.
#include <vector>

#include "common.cuh"

extern "C" int64_t b2asr_edit_distance(const int32_t* a, int64_t na, const int32_t* b, int64_t nb) {
  if (na < 0 || nb < 0 || (na > 0 && !a) || (nb > 0 && !b)) return -1;
  if (na < nb) { const int32_t* t = a; a = b; b = t; const int64_t n = na; na = nb; nb = n; }   // b is the shorter one
  std::vector<int64_t> row((size_t)nb + 1);
  for (int64_t j = 0; j <= nb; ++j) row[(size_t)j] = j;
  for (int64_t i = 1; i <= na; ++i) {
    int64_t diag = row[0];
    row[0] = i;
    const int32_t x = a[i - 1];
    for (int64_t j = 1; j <= nb; ++j) {
      const int64_t up = row[(size_t)j];
      int64_t best = diag + (x != b[j - 1]);
      if (up + 1 < best) best = up + 1;
      if (row[(size_t)j - 1] + 1 < best) best = row[(size_t)j - 1] + 1;
      row[(size_t)j] = best;
      diag = up;
    }
  }
  return row[(size_t)nb];
}
