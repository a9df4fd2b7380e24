/*
 * cub_sort.cu -- cub::DeviceRadixSort wrapper.  Used only by rt_sort_pairs_cub, the
 * cross-check tests/test_gpu_bvh.py runs against the builder's own radix sort
 * (bvh_build.cu); the product path never calls it.
 */
#include <cub/device/device_radix_sort.cuh>
#include <cuda_runtime.h>

extern "C" cudaError_t rt_cub_sort_pairs(unsigned int* keys, unsigned int* vals, int n, cudaStream_t stream) {
    unsigned int *ko = nullptr, *vo = nullptr;
    void* tmp = nullptr;
    size_t tmp_bytes = 0;
    cudaError_t e = cudaMalloc((void**)&ko, sizeof(unsigned int) * (size_t)n);
    if (e != cudaSuccess) return e;
    e = cudaMalloc((void**)&vo, sizeof(unsigned int) * (size_t)n);
    if (e != cudaSuccess) { cudaFree(ko); return e; }
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, ko, vals, vo, n, 0, 32, stream);
    e = cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1);
    if (e == cudaSuccess) {
        e = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, ko, vals, vo, n, 0, 32, stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(keys, ko, sizeof(unsigned int) * (size_t)n, cudaMemcpyDeviceToDevice, stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(vals, vo, sizeof(unsigned int) * (size_t)n, cudaMemcpyDeviceToDevice, stream);
        cudaStreamSynchronize(stream);
    }
    cudaFree(tmp); cudaFree(ko); cudaFree(vo);
    return e;
}
