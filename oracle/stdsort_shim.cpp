// oracle/stdsort_shim.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// libstdc++'s std::sort with the reference's comparator (`a.response > b.response`,
// trackFeatures.cpp:404-406) over the oracle's keypoint records.  The permutation std::sort
// produces depends only on the element count, the input order and the comparison results, so
// with this installed as oracle/lk_oracle.c's sort hook (olk_set_sort_hook) the restatement
// orders tied responses exactly as the reference objects built by the same toolchain do.
// Used by tests/test_ref_lk.py to separate "the restatement misreads the reference" from
// "the tie order of an unstable sort differs" (DESIGN.md section 2).
#include <algorithm>
#include <cstddef>

struct olk_kp { int x, y, resp; };

static bool by_response(const olk_kp &a, const olk_kp &b) { return (float)a.resp > (float)b.resp; }

extern "C" void olk_std_sort(olk_kp *kps, size_t n, int /*W*/) { std::sort(kps, kps + n, by_response); }
