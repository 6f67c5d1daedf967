// py::Tensor and py::CudaSync of the reference (include/madrona/python.hpp:21-73, src/python/utils.cpp:15-44):
// what an executor hands to a Python trainer. Tensor = (pointer, element type, <= 16 dimensions, gpu id or
// -1 for host memory): a non-owning description of an exported column; CudaSync = the stream-ordered hand-off
// (the reference waits an external semaphore signalled by its renderer on the trainer's stream; here the
// executor records a CUDA event after the last step it enqueued, and wait() makes the trainer's stream
// wait for it -- no host synchronisation).
#pragma once
#include <array>
#include <cstdint>

#include "../../madrona_b200.h"

namespace madrona {
namespace py {

class Tensor final {
public:
    enum class ElementType { UInt8, Int8, Int16, Int32, Int64, Float16, Float32 };
    static inline constexpr int64_t maxDimensions = 16;

    Tensor(void *dev_ptr, ElementType type, const int64_t *dimensions, int64_t num_dimensions, int gpu_id /* -1: host */)
        : dev_ptr_(dev_ptr), type_(type), gpu_id_(gpu_id), num_dimensions_(num_dimensions)
    {
        for (int64_t i = 0; i < num_dimensions && i < maxDimensions; i++) dimensions_[(size_t)i] = dimensions[i];
    }
    inline void *devicePtr() const { return dev_ptr_; }
    inline ElementType type() const { return type_; }
    inline bool isOnGPU() const { return gpu_id_ != -1; }
    inline int gpuID() const { return gpu_id_; }
    inline int64_t numDims() const { return num_dimensions_; }
    inline const int64_t *dims() const { return dimensions_.data(); }
    static inline int64_t elementBytes(ElementType t)
    {
        switch (t) {
        case ElementType::UInt8: case ElementType::Int8: return 1;
        case ElementType::Int16: case ElementType::Float16: return 2;
        case ElementType::Int32: case ElementType::Float32: return 4;
        case ElementType::Int64: return 8;
        }
        return 0;
    }

private:
    void *dev_ptr_;
    ElementType type_;
    int gpu_id_;
    int64_t num_dimensions_;
    std::array<int64_t, maxDimensions> dimensions_ {};
};

class CudaSync final {
public:
    explicit CudaSync(void *done_event) : event_(done_event) {}
    // `strm` = the raw cudaStream_t of the consumer (torch.cuda.current_stream().cuda_stream)
    inline void wait(uint64_t strm) { mdr_stream_wait_event((void *)strm, event_); }

private:
    void *event_;
};

}  // namespace py
}  // namespace madrona
