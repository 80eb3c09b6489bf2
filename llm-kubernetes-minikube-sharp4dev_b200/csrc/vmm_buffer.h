// vmm_buffer.h -- append-only device buffers whose base address never moves.
//
// The reference's store is an unbounded `List<RagChunk>` (Helpers/RagIndex.cs:8).  Its device
// counterpart reserves virtual address space for the whole capacity up front (free) and maps
// physical HBM behind it chunk by chunk as rows arrive (CUDA virtual memory management), so
// concurrent searches keep valid pointers while the store grows and a mostly-empty store costs
// almost nothing of the 180 GB.  Driver entry points are fetched through
// cudaGetDriverEntryPoint, so libragstore.so has no link-time dependency on libcuda.so and
// still loads (and fails cleanly) on a machine without a GPU.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <vector>

namespace rag {

struct VmmApi {
    decltype(&cuMemAddressReserve) addressReserve = nullptr;
    decltype(&cuMemAddressFree) addressFree = nullptr;
    decltype(&cuMemCreate) create = nullptr;
    decltype(&cuMemRelease) release = nullptr;
    decltype(&cuMemMap) map = nullptr;
    decltype(&cuMemUnmap) unmap = nullptr;
    decltype(&cuMemSetAccess) setAccess = nullptr;
    decltype(&cuMemGetAllocationGranularity) granularity = nullptr;
    bool ok = false;

    static const VmmApi& get() {
        static VmmApi api = [] {
            VmmApi a;
            auto load = [](const char* name, void** fn) {
                cudaDriverEntryPointQueryResult q;
                return cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) == cudaSuccess &&
                       q == cudaDriverEntryPointSuccess && *fn != nullptr;
            };
            a.ok = load("cuMemAddressReserve", reinterpret_cast<void**>(&a.addressReserve)) &&
                   load("cuMemAddressFree", reinterpret_cast<void**>(&a.addressFree)) &&
                   load("cuMemCreate", reinterpret_cast<void**>(&a.create)) &&
                   load("cuMemRelease", reinterpret_cast<void**>(&a.release)) &&
                   load("cuMemMap", reinterpret_cast<void**>(&a.map)) &&
                   load("cuMemUnmap", reinterpret_cast<void**>(&a.unmap)) &&
                   load("cuMemSetAccess", reinterpret_cast<void**>(&a.setAccess)) &&
                   load("cuMemGetAllocationGranularity", reinterpret_cast<void**>(&a.granularity));
            cudaGetLastError();
            return a;
        }();
        return api;
    }
};

class VmmBuffer {
public:
    // Reserve address space for max_bytes on `device`.  Returns cudaError_t-compatible int (0 = ok).
    int reserve(size_t max_bytes, int device) {
        device_ = device;
        const VmmApi& api = VmmApi::get();
        if (api.ok) {
            CUmemAllocationProp prop = {};
            prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
            prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
            prop.location.id = device;
            size_t g = 0;
            if (api.granularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && g > 0) {
                gran_ = g;
                reserved_ = (std::max<size_t>(max_bytes, 1) + gran_ - 1) / gran_ * gran_;
                CUdeviceptr p = 0;
                if (api.addressReserve(&p, reserved_, 0, 0, 0) == CUDA_SUCCESS) {
                    base_ = p;
                    vmm_ = true;
                    return 0;
                }
            }
        }
        // no VMM (very old driver): one plain allocation of the whole capacity
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, std::max<size_t>(max_bytes, 256));
        if (e != cudaSuccess) return static_cast<int>(e);
        base_ = reinterpret_cast<CUdeviceptr>(p);
        reserved_ = mapped_ = std::max<size_t>(max_bytes, 256);
        vmm_ = false;
        return 0;
    }

    // Make [0, bytes) usable.  Newly mapped memory is zero-filled on `stream` when zero==true.
    int ensure(size_t bytes, bool zero, cudaStream_t stream) {
        if (bytes <= mapped_) return 0;
        if (bytes > reserved_) return static_cast<int>(cudaErrorMemoryAllocation);
        if (!vmm_) return 0;
        const VmmApi& api = VmmApi::get();
        // grow geometrically (at least the request; doubling, capped at +16 GiB) in granularity units, so that a
        // store fed by many appends maps memory O(log) times; if the generous amount is not available, map the need
        const size_t need = (bytes - mapped_ + gran_ - 1) / gran_ * gran_;
        size_t want = std::max(bytes - mapped_, std::min<size_t>(mapped_, size_t(16) << 30));
        want = (want + gran_ - 1) / gran_ * gran_;
        want = std::min(want, reserved_ - mapped_);
        CUmemAllocationProp prop = {};
        prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        prop.location.id = device_;
        CUmemGenericAllocationHandle h;
        if (api.create(&h, want, &prop, 0) != CUDA_SUCCESS) {
            want = std::min(need, reserved_ - mapped_);
            if (api.create(&h, want, &prop, 0) != CUDA_SUCCESS) return static_cast<int>(cudaErrorMemoryAllocation);
        }
        if (api.map(base_ + mapped_, want, 0, h, 0) != CUDA_SUCCESS) {
            api.release(h);
            return static_cast<int>(cudaErrorMemoryAllocation);
        }
        CUmemAccessDesc acc = {};
        acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc.location.id = device_;
        acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        if (api.setAccess(base_ + mapped_, want, &acc, 1) != CUDA_SUCCESS) {
            api.unmap(base_ + mapped_, want);
            api.release(h);
            return static_cast<int>(cudaErrorMemoryAllocation);
        }
        handles_.push_back(h);
        sizes_.push_back(want);
        if (zero) {
            cudaError_t e = cudaMemsetAsync(reinterpret_cast<void*>(base_ + mapped_), 0, want, stream);
            if (e != cudaSuccess) return static_cast<int>(e);
        }
        mapped_ += want;
        return 0;
    }

    void release() {
        if (!base_) return;
        if (vmm_) {
            const VmmApi& api = VmmApi::get();
            size_t off = 0;
            for (size_t i = 0; i < handles_.size(); ++i) {
                api.unmap(base_ + off, sizes_[i]);
                api.release(handles_[i]);
                off += sizes_[i];
            }
            api.addressFree(base_, reserved_);
        } else {
            cudaFree(reinterpret_cast<void*>(base_));
        }
        handles_.clear();
        sizes_.clear();
        base_ = 0;
        reserved_ = mapped_ = 0;
    }

    template <typename T>
    T* ptr() const {
        return reinterpret_cast<T*>(base_);
    }
    size_t mapped() const { return mapped_; }
    size_t reserved() const { return reserved_; }
    bool uses_vmm() const { return vmm_; }

private:
    CUdeviceptr base_ = 0;
    size_t reserved_ = 0, mapped_ = 0, gran_ = 0;
    int device_ = 0;
    bool vmm_ = false;
    std::vector<CUmemGenericAllocationHandle> handles_;
    std::vector<size_t> sizes_;
};

}  // namespace rag
