// slc_prims.cu -- device-wide building blocks: prefix scans and the stable LSD radix sort
// that transpose and triplet assembly are built on.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "slc_common.cuh"

namespace slc {

template <class U>
static int exclusive_scan_impl(const U *in, U *out, int64_t n) {
  if (n <= 0) return SLC_OK;
  size_t tmp_bytes = 0;
  SLC_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, in, out, n, cur_stream()));
  DevBuf<char> tmp;
  SLC_TRY(tmp.alloc(tmp_bytes));
  SLC_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, in, out, n, cur_stream()));
  g_launches.fetch_add(2, std::memory_order_relaxed);
  return SLC_OK;
}

int exclusive_scan_i32(const int32_t *in, int32_t *out, int64_t n) {
  return exclusive_scan_impl<int32_t>(in, out, n);
}
int exclusive_scan_i64(const int64_t *in, int64_t *out, int64_t n) {
  return exclusive_scan_impl<int64_t>(in, out, n);
}

struct MaxOp {
  __host__ __device__ __forceinline__ int32_t operator()(int32_t a, int32_t b) const {
    return a > b ? a : b;
  }
};

int inclusive_max_scan_i32(const int32_t *in, int32_t *out, int64_t n) {
  if (n <= 0) return SLC_OK;
  size_t tmp_bytes = 0;
  SLC_CUDA(cub::DeviceScan::InclusiveScan(nullptr, tmp_bytes, in, out, MaxOp(), n, cur_stream()));
  DevBuf<char> tmp;
  SLC_TRY(tmp.alloc(tmp_bytes));
  SLC_CUDA(cub::DeviceScan::InclusiveScan(tmp.p, tmp_bytes, in, out, MaxOp(), n, cur_stream()));
  g_launches.fetch_add(2, std::memory_order_relaxed);
  return SLC_OK;
}

template <class K>
static int sort_pairs_impl(K *keys_in, K *keys_out, uint32_t *vals_in, uint32_t *vals_out, int64_t n,
                           int end_bit) {
  if (n <= 0) return SLC_OK;
  if (end_bit < 1) end_bit = 1;
  size_t tmp_bytes = 0;
  SLC_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys_in, keys_out, vals_in, vals_out, n,
                                           0, end_bit, cur_stream()));
  DevBuf<char> tmp;
  SLC_TRY(tmp.alloc(tmp_bytes));
  SLC_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, keys_in, keys_out, vals_in, vals_out, n, 0,
                                           end_bit, cur_stream()));
  g_launches.fetch_add(1 + (end_bit + 7) / 8, std::memory_order_relaxed);
  return SLC_OK;
}

int sort_pairs_u32(uint32_t *keys_in, uint32_t *keys_out, uint32_t *vals_in, uint32_t *vals_out,
                   int64_t n, int end_bit) {
  return sort_pairs_impl<uint32_t>(keys_in, keys_out, vals_in, vals_out, n, end_bit);
}
int sort_pairs_u64(uint64_t *keys_in, uint64_t *keys_out, uint32_t *vals_in, uint32_t *vals_out,
                   int64_t n, int end_bit) {
  return sort_pairs_impl<uint64_t>(keys_in, keys_out, vals_in, vals_out, n, end_bit);
}

}  // namespace slc
