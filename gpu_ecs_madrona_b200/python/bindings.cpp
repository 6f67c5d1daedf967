// The `madrona_python` module of the reference (src/python/bindings.cpp:78-128): class Tensor with
// Tensor(torch_tensor) and .to_torch(), class CudaSync with .wait(stream). The reference binds with nanobind
// (vendored under its external/); this image has pybind11 and no nanobind, so the same module surface is
// written against pybind11, with the DLPack capsule built by hand (DLPack 0.8 `dltensor` protocol:
// __dlpack__ / __dlpack_device__, so torch.from_dlpack(t), cupy.from_dlpack(t), jax ... all take it).
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cstdint>
#include <stdexcept>
#include <vector>

#include "../../include/madrona/b200/python.hpp"

namespace pyb = pybind11;
using madrona::py::CudaSync;
using madrona::py::Tensor;
using ET = Tensor::ElementType;

namespace {

// ---- DLPack (dlpack.h, v0.8) ----
enum { kDLCPU = 1, kDLCUDA = 2 };
enum { kDLInt = 0, kDLUInt = 1, kDLFloat = 2 };
struct DLDevice { int32_t device_type; int32_t device_id; };
struct DLDataType { uint8_t code; uint8_t bits; uint16_t lanes; };
struct DLTensor {
    void *data; DLDevice device; int32_t ndim; DLDataType dtype;
    int64_t *shape; int64_t *strides; uint64_t byte_offset;
};
struct DLManagedTensor { DLTensor dl_tensor; void *manager_ctx; void (*deleter)(DLManagedTensor *); };

struct Managed {
    DLManagedTensor m;
    int64_t shape[Tensor::maxDimensions];
    PyObject *owner;             // keeps the Python Tensor (and through it nothing else) alive
};

DLDataType toDLPackType(ET t)
{
    switch (t) {
    case ET::UInt8: return { kDLUInt, 8, 1 };
    case ET::Int8: return { kDLInt, 8, 1 };
    case ET::Int16: return { kDLInt, 16, 1 };
    case ET::Int32: return { kDLInt, 32, 1 };
    case ET::Int64: return { kDLInt, 64, 1 };
    case ET::Float16: return { kDLFloat, 16, 1 };
    case ET::Float32: return { kDLFloat, 32, 1 };
    }
    throw std::invalid_argument("Tensor: invalid element type");
}

ET fromDLPackType(DLDataType d)
{
    if (d.lanes == 1) {
        if (d.code == kDLInt && d.bits == 8) return ET::Int8;
        if (d.code == kDLInt && d.bits == 16) return ET::Int16;
        if (d.code == kDLInt && d.bits == 32) return ET::Int32;
        if (d.code == kDLInt && d.bits == 64) return ET::Int64;
        if (d.code == kDLUInt && d.bits == 8) return ET::UInt8;
        if (d.code == kDLFloat && d.bits == 16) return ET::Float16;
        if (d.code == kDLFloat && d.bits == 32) return ET::Float32;
    }
    throw std::invalid_argument("Tensor: Invalid tensor dtype");     // the reference aborts here
}

void capsuleDestructor(PyObject *cap)
{
    // a consumer renames the capsule to "used_dltensor" and takes over the deleter
    if (!PyCapsule_IsValid(cap, "dltensor")) return;
    auto *m = (DLManagedTensor *)PyCapsule_GetPointer(cap, "dltensor");
    if (m && m->deleter) m->deleter(m);
}

pyb::object makeCapsule(const Tensor &t, pyb::handle owner)
{
    Managed *mg = new Managed;
    for (int64_t i = 0; i < t.numDims(); i++) mg->shape[i] = t.dims()[i];
    mg->owner = owner.ptr();
    Py_XINCREF(mg->owner);
    DLTensor &d = mg->m.dl_tensor;
    d.data = t.devicePtr();
    d.device = t.isOnGPU() ? DLDevice { kDLCUDA, t.gpuID() } : DLDevice { kDLCPU, 0 };
    d.ndim = (int32_t)t.numDims();
    d.dtype = toDLPackType(t.type());
    d.shape = mg->shape;
    d.strides = nullptr;          // compact row-major
    d.byte_offset = 0;
    mg->m.manager_ctx = mg;
    mg->m.deleter = [](DLManagedTensor *m) {
        Managed *self = (Managed *)m->manager_ctx;
        if (self->owner) {
            PyGILState_STATE g = PyGILState_Ensure();
            Py_DECREF(self->owner);
            PyGILState_Release(g);
        }
        delete self;
    };
    return pyb::reinterpret_steal<pyb::object>(PyCapsule_New(&mg->m, "dltensor", capsuleDestructor));
}

Tensor fromDLPackObject(pyb::object src)
{
    // Tensor(torch_tensor): import through the producer's own __dlpack__
    pyb::object cap = src.attr("__dlpack__")();
    if (!PyCapsule_IsValid(cap.ptr(), "dltensor")) throw std::invalid_argument("Tensor: not a DLPack capsule");
    auto *m = (DLManagedTensor *)PyCapsule_GetPointer(cap.ptr(), "dltensor");
    const DLTensor &d = m->dl_tensor;
    if (d.ndim > Tensor::maxDimensions) throw std::invalid_argument("Cannot construct Tensor with more than 16 dimensions");
    int gpu = -1;
    if (d.device.device_type == kDLCUDA) gpu = d.device.device_id;
    else if (d.device.device_type != kDLCPU) throw std::invalid_argument("madrona::Tensor: failed to import unknown tensor type");
    if (d.strides) {
        int64_t expect = 1;
        for (int i = d.ndim - 1; i >= 0; i--) {
            if (d.shape[i] != 1 && d.strides[i] != expect) throw std::invalid_argument("Tensor: source is not contiguous");
            expect *= d.shape[i];
        }
    }
    Tensor t((char *)d.data + d.byte_offset, fromDLPackType(d.dtype), d.shape, d.ndim, gpu);
    // the view does not own the memory (as in the reference): the capsule is released unconsumed
    return t;
}

}  // namespace

PYBIND11_MODULE(madrona_python, m)
{
    m.doc() = "Tensor / CudaSync hand-off objects of the B200 world-step executors (madrona_python of the reference)";
    pyb::enum_<ET>(m, "ElementType")
        .value("UInt8", ET::UInt8).value("Int8", ET::Int8).value("Int16", ET::Int16).value("Int32", ET::Int32)
        .value("Int64", ET::Int64).value("Float16", ET::Float16).value("Float32", ET::Float32);

    pyb::class_<Tensor>(m, "Tensor")
        .def(pyb::init(&fromDLPackObject), pyb::arg("torch_tensor"))
        .def_static("from_pointer", [](uint64_t ptr, ET type, std::vector<int64_t> dims, int gpu_id) {
            if ((int64_t)dims.size() > Tensor::maxDimensions) throw std::invalid_argument("Cannot construct Tensor with more than 16 dimensions");
            return Tensor((void *)ptr, type, dims.data(), (int64_t)dims.size(), gpu_id);
        }, pyb::arg("ptr"), pyb::arg("type"), pyb::arg("dims"), pyb::arg("gpu_id") = -1,
           "Tensor(void *, ElementType, dims, gpu id) of the C++ side: what Executor::getExported(slot) is wrapped in")
        .def("__dlpack__", [](pyb::object self, pyb::object /*stream*/) { return makeCapsule(self.cast<const Tensor &>(), self); },
             pyb::arg("stream") = pyb::none())
        .def("__dlpack_device__", [](const Tensor &t) {
            return t.isOnGPU() ? pyb::make_tuple((int)kDLCUDA, t.gpuID()) : pyb::make_tuple((int)kDLCPU, 0);
        })
        .def("to_torch", [](pyb::object self) {
            pyb::object from_dlpack = pyb::module_::import("torch.utils.dlpack").attr("from_dlpack");
            return from_dlpack(makeCapsule(self.cast<const Tensor &>(), self));
        })
        .def_property_readonly("ptr", [](const Tensor &t) { return (uint64_t)t.devicePtr(); })
        .def_property_readonly("type", &Tensor::type)
        .def_property_readonly("gpu_id", &Tensor::gpuID)
        .def_property_readonly("is_on_gpu", &Tensor::isOnGPU)
        .def_property_readonly("dims", [](const Tensor &t) { return std::vector<int64_t>(t.dims(), t.dims() + t.numDims()); });

    pyb::class_<CudaSync>(m, "CudaSync")
        .def(pyb::init([](uint64_t done_event) { return CudaSync((void *)done_event); }), pyb::arg("done_event"))
        .def("wait", &CudaSync::wait, pyb::arg("stream"));
}
