// Growable device arrays over CUDA virtual memory management (see mmg_vm.h).
#include "mmg_vm.h"

#include <mutex>

namespace mmg
{
namespace
{
struct VmApi
{
    CUresult (*reserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long)                       = nullptr;
    CUresult (*addr_free)(CUdeviceptr, size_t)                                                                = nullptr;
    CUresult (*create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*release)(CUmemGenericAllocationHandle)                                                         = nullptr;
    CUresult (*map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long)            = nullptr;
    CUresult (*unmap)(CUdeviceptr, size_t)                                                                    = nullptr;
    CUresult (*set_access)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t)                               = nullptr;
    CUresult (*granularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags)            = nullptr;
    bool ok = false;
};
template <typename F>
bool entry(const char* name, F& fn)
{
    void*                           p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p)
    {
        cudaGetLastError();
        return false;
    }
    fn = reinterpret_cast<F>(p);
    return true;
}
VmApi& api()
{
    static VmApi          a;
    static std::once_flag once;
    std::call_once(once, [] {
        a.ok = entry("cuMemAddressReserve", a.reserve) && entry("cuMemAddressFree", a.addr_free) && entry("cuMemCreate", a.create) &&
               entry("cuMemRelease", a.release) && entry("cuMemMap", a.map) && entry("cuMemUnmap", a.unmap) && entry("cuMemSetAccess", a.set_access) &&
               entry("cuMemGetAllocationGranularity", a.granularity);
    });
    return a;
}
CUmemAllocationProp prop_for(int device)
{
    CUmemAllocationProp p = {};
    p.type                = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.location.type       = CU_MEM_LOCATION_TYPE_DEVICE;
    p.location.id         = device;
    return p;
}
}  // namespace

bool vm_available() { return api().ok; }

int vm_reserve(VmArray& a, size_t max_bytes, int device)
{
    VmApi& A = api();
    if (!A.ok) return -1;
    const CUmemAllocationProp p = prop_for(device);
    size_t                    g = 0;
    if (A.granularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || g == 0) return -1;
    a.gran     = g;
    a.device   = device;
    a.reserved = (max_bytes + g - 1) / g * g;
    if (A.reserve(&a.base, a.reserved, 0, 0, 0) != CUDA_SUCCESS)
    {
        a.base = 0, a.reserved = 0;
        return -1;
    }
    return 0;
}

int vm_ensure(VmArray& a, size_t bytes)
{
    if (bytes <= a.mapped) return 0;
    VmApi& A = api();
    if (!a.base || bytes > a.reserved) return -1;
    const size_t                 want = (bytes + a.gran - 1) / a.gran * a.gran, add = want - a.mapped;
    const CUmemAllocationProp    p    = prop_for(a.device);
    CUmemGenericAllocationHandle h;
    if (A.create(&h, add, &p, 0) != CUDA_SUCCESS) return -1;
    if (A.map(a.base + a.mapped, add, 0, h, 0) != CUDA_SUCCESS)
    {
        A.release(h);
        return -1;
    }
    CUmemAccessDesc d = {};
    d.location        = p.location;
    d.flags           = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    if (A.set_access(a.base + a.mapped, add, &d, 1) != CUDA_SUCCESS)
    {
        A.unmap(a.base + a.mapped, add);
        A.release(h);
        return -1;
    }
    a.handles.push_back(h);
    a.sizes.push_back(add);
    a.mapped = want;
    return 0;
}

void vm_release(VmArray& a)
{
    VmApi& A = api();
    if (!a.base) return;
    size_t off = 0;
    for (size_t i = 0; i < a.handles.size(); i++)
    {
        A.unmap(a.base + off, a.sizes[i]);
        A.release(a.handles[i]);
        off += a.sizes[i];
    }
    A.addr_free(a.base, a.reserved);
    a = VmArray();
}
}  // namespace mmg
