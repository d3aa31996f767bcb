// Gradient-bucket assignment by size.
//
// Contract reproduced (what the reference gets from torch DDP's constructor, see
// launch_dist.py:61 / example_launch.py:29): tensors are visited in the given order; tensors
// of different "kind" (dtype/device key) never share a bucket; every kind keeps its own open
// bucket and its own cursor into the list of size limits; a bucket is closed right AFTER the
// tensor that makes it reach the limit (soft cap); leftovers are flushed at the end; buckets
// are finally ordered by their smallest tensor index.
#include <algorithm>
#include <cstdint>
#include <map>
#include <vector>

extern "C" {

// sizes[i]  : bytes of tensor i          kinds[i] : grouping key of tensor i
// limits[j] : byte caps; the last one repeats forever
// bucket_of[i] (out) : bucket index of tensor i.  Returns the number of buckets.
int pdl_bucket_assign(const int64_t* sizes, const int32_t* kinds, int n, const int64_t* limits, int nlimits, int32_t* bucket_of) {
  struct Open {
    std::vector<int> idx;
    int64_t bytes = 0;
    int limit_pos = 0;
  };
  std::map<int32_t, Open> open;
  std::vector<std::vector<int>> done;
  for (int i = 0; i < n; ++i) {
    Open& b = open[kinds ? kinds[i] : 0];
    b.idx.push_back(i);
    b.bytes += sizes[i];
    int64_t cap = limits[std::min(b.limit_pos, nlimits - 1)];
    if (b.bytes >= cap) {
      done.push_back(std::move(b.idx));
      b.idx.clear();
      b.bytes = 0;
      if (b.limit_pos < nlimits - 1) ++b.limit_pos;
    }
  }
  for (auto& kv : open)
    if (!kv.second.idx.empty()) done.push_back(std::move(kv.second.idx));
  std::sort(done.begin(), done.end(), [](const std::vector<int>& a, const std::vector<int>& b) {
    return *std::min_element(a.begin(), a.end()) < *std::min_element(b.begin(), b.end());
  });
  for (size_t k = 0; k < done.size(); ++k)
    for (int i : done[k]) bucket_of[i] = static_cast<int32_t>(k);
  return static_cast<int>(done.size());
}

}  // extern "C"
