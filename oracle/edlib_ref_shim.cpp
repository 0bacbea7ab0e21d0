// TEST INFRASTRUCTURE: C entry point over the reference's vendored edlib (compiled from /root/reference by oracle/Makefile
// into oracle/_ref/): global (NW) edit distance, what edlib.align(a, b)["editDistance"] returns (remove_redundancy.py:75-81).
#include <cstddef>

#include "edlib.h"
extern "C" int edlib_ref_distance(const char* a, int na, const char* b, int nb) {
  EdlibAlignResult r = edlibAlign(a, na, b, nb, edlibNewAlignConfig(-1, EDLIB_MODE_NW, EDLIB_TASK_DISTANCE, NULL, 0));
  const int d = r.status == EDLIB_STATUS_OK ? r.editDistance : -2;
  edlibFreeAlignResult(r);
  return d;
}
