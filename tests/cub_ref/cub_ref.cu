// TEST-ONLY reference (never linked into libsplat_b200.so): cub::DeviceRadixSort / cub::DeviceScan behind a
// minimal C ABI, as an independent checker of csrc/radix_sort.cu at sizes torch.sort is slow for, and as the
// library baseline scripts/bench_sort.py times the hand-written sort against.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <stdint.h>

#define API extern "C" __attribute__((visibility("default")))

API size_t cubref_sort64_temp(int count) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, (const int*)nullptr,
                                  (int*)nullptr, count, 0, 64);
  return bytes;
}
API int cubref_sort64(int count, int begin_bit, int end_bit, const uint64_t* kin, const int* vin, uint64_t* kout,
                      int* vout, void* temp, size_t temp_bytes, void* stream) {
  return (int)cub::DeviceRadixSort::SortPairs(temp, temp_bytes, kin, kout, vin, vout, count, begin_bit, end_bit,
                                              (cudaStream_t)stream);
}
API size_t cubref_sort32_temp(int count) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const int*)nullptr,
                                  (int*)nullptr, count, 0, 32);
  return bytes;
}
API int cubref_sort32(int count, int begin_bit, int end_bit, const uint32_t* kin, const int* vin, uint32_t* kout,
                      int* vout, void* temp, size_t temp_bytes, void* stream) {
  return (int)cub::DeviceRadixSort::SortPairs(temp, temp_bytes, kin, kout, vin, vout, count, begin_bit, end_bit,
                                              (cudaStream_t)stream);
}
API size_t cubref_scan_temp(int count) {
  size_t bytes = 0;
  cub::DeviceScan::InclusiveSum(nullptr, bytes, (const int*)nullptr, (int*)nullptr, count);
  return bytes;
}
API int cubref_scan(int count, const int* in, int* out, void* temp, size_t temp_bytes, void* stream) {
  return (int)cub::DeviceScan::InclusiveSum(temp, temp_bytes, in, out, count, (cudaStream_t)stream);
}
