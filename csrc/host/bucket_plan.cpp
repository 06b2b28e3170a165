// b200flow — DDP gradient-bucket planner (host C++).
//
// torch's C++ `Reducer` assigns parameters to ~25 MB buckets in reverse registration
// order and launches one all-reduce per bucket as soon as its gradients are ready
// (SURVEY.md §2F "DP", §2H C1; Horovod's fusion buffer plays the same role for the
// Mask R-CNN workloads, C2).  With gradients living in one flat buffer a bucket is a
// contiguous slice, so the planner only has to choose cut points: walk the slots
// from the END of the buffer (the order backward produces gradients in) and close a
// bucket when it reaches the byte cap; the first bucket is kept small so the first
// collective starts early.
#include <cstdint>

extern "C" {

// offsets/numels: per-parameter slot (elements) in registration order.
// Writes up to max_buckets rows of {first_slot, last_slot_exclusive, elem_begin, elem_end}
// (buckets ordered as backward fills them: last parameters first). Returns bucket count.
int b2f_plan_buckets(const int64_t* offsets, const int64_t* numels, int nslots,
                     int64_t total_elems, int64_t elem_bytes, int64_t cap_bytes,
                     int64_t first_cap_bytes, int64_t* out, int max_buckets) {
  int nb = 0;
  int hi_slot = nslots;
  int64_t hi_elem = total_elems;
  int64_t acc = 0;
  for (int s = nslots - 1; s >= 0; --s) {
    acc += numels[s] * elem_bytes;
    const int64_t cap = nb == 0 ? first_cap_bytes : cap_bytes;
    if (acc >= cap || s == 0) {
      if (nb >= max_buckets) return -1;
      const int64_t lo_elem = s == 0 ? 0 : offsets[s];
      out[4 * nb + 0] = s;
      out[4 * nb + 1] = hi_slot;
      out[4 * nb + 2] = lo_elem;
      out[4 * nb + 3] = hi_elem;
      ++nb;
      hi_slot = s;
      hi_elem = lo_elem;
      acc = 0;
    }
  }
  return nb;
}

}  // extern "C"
