// Device arrays that grow in place: a large virtual range is reserved once and physical memory is mapped behind it as
// the array grows (CUDA virtual memory management), so growth never copies the old contents, never frees, and the
// base pointer never changes. The driver entry points are looked up at run time (no link against libcuda).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstddef>
#include <vector>

namespace mmg
{
struct VmArray
{
    CUdeviceptr                               base = 0;
    size_t                                    reserved = 0, mapped = 0, gran = 0;
    int                                       device = 0;
    std::vector<CUmemGenericAllocationHandle> handles;
    std::vector<size_t>                       sizes;
};
// false when the driver lacks the API (the caller then uses plain cudaMalloc growth)
bool vm_available();
// reserves `max_bytes` of address space (no physical memory); 0 or -1
int vm_reserve(VmArray& a, size_t max_bytes, int device);
// makes [0, bytes) usable; new memory is NOT initialised. Returns 0, or -1 on failure (out of memory / past the range)
int vm_ensure(VmArray& a, size_t bytes);
void vm_release(VmArray& a);
}  // namespace mmg
