// cub_sort.cu — yardstick only (never linked into the product): cub::DeviceRadixSort::SortPairs (the CUDA toolkit's onesweep) on
// the bench's standalone sort workloads, to place the hand-written K4 against the library on the same B200.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/ubench/cub_sort tools/ubench/cub_sort.cu
#include <cub/device/device_radix_sort.cuh>
#include <cstdio>
#include <cstdint>
#include <vector>

static uint64_t splitmix(uint64_t& s) { uint64_t z = (s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }

template <typename K>
void run(int logn, int reps) {
    const size_t n = 1ull << logn;
    std::vector<K> h(n); uint64_t s = 1234 + logn;
    for (auto& k : h) k = (K)splitmix(s);
    K *ki, *ko; uint32_t *vi, *vo;
    cudaMalloc(&ki, n * sizeof(K)); cudaMalloc(&ko, n * sizeof(K)); cudaMalloc(&vi, n * 4); cudaMalloc(&vo, n * 4);
    cudaMemcpy(ki, h.data(), n * sizeof(K), cudaMemcpyHostToDevice); cudaMemset(vi, 0, n * 4);
    size_t tb = 0; cub::DeviceRadixSort::SortPairs(nullptr, tb, ki, ko, vi, vo, n);
    void* tmp; cudaMalloc(&tmp, tb);
    for (int i = 0; i < 3; i++) cub::DeviceRadixSort::SortPairs(tmp, tb, ki, ko, vi, vo, n);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int i = 0; i < reps; i++) cub::DeviceRadixSort::SortPairs(tmp, tb, ki, ko, vi, vo, n);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= reps;
    printf("cub SortPairs u%zu+u32 n=2^%d: %.3f ms  %.2f Gkeys/s\n", sizeof(K) * 8, logn, ms, n / ms / 1e6);
    cudaFree(ki); cudaFree(ko); cudaFree(vi); cudaFree(vo); cudaFree(tmp);
}
int main() {
    for (int l : {20, 22, 24, 26}) run<uint64_t>(l, 10);
    for (int l : {20, 24}) run<uint32_t>(l, 10);
    return 0;
}
