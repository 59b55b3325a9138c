// array.hpp — the two containers that cross the class-API boundary, with the reference's interface:
//   Array<T,N>      host, column-major, 1-based indices            (reference include/Array.h:94-297)
//   Array_gpu<T,N>  device, owning or a view on caller memory      (reference include/Array.h:355-650)
// Same member names, argument meaning and error behaviour, so code written against the reference
// (load_and_init_gas_optics, src_test/Radiation_solver.cu:93-243; MicroHH call sites that subset or convert)
// compiles unchanged.  Differences, all in the direction of fewer hidden synchronisations:
//   * device memory is stream-ordered (cudaMallocAsync / cudaFreeAsync on rrtmgp_b200::stream());
//   * fill() and subset() are kernels of librrtmgp_b200.so enqueued on that stream (the reference launches on
//     the legacy default stream); D2D copies are cudaMemcpyAsync on it;
//   * operator()(indices) on a device array still has to block (it returns a host value): it synchronises the
//     stream, not the device.
// This header is plain C++ (g++ or nvcc): the kernels live behind the C ABI (include/rrtmgp_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <algorithm>
#include <array>
#include <fstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>
#include "../rrtmgp_b200.h"

namespace rrtmgp_b200 {
inline cudaStream_t& stream() { static cudaStream_t s = nullptr; return s; }
inline void check(int rc)
{
    if (rc == RRTMGP_B200_OK) return;
    const std::string msg = rrtmgp_b200_last_error();
    if (rc == RRTMGP_B200_E_RANGE) throw std::range_error(msg);      // src/Gas_optics_rrtmgp.cpp:812-827
    throw std::runtime_error(msg);
}
inline void cuda_check(cudaError_t e) { if (e != cudaSuccess) throw std::runtime_error(cudaGetErrorString(e)); }

// Shape bookkeeping shared by both containers: extents, column-major strides, index offsets.
// Element (i1..iN) lives at sum_k (i_k - offset_k - 1) * stride_k   (include/Array.h:52-63).
template<int N> struct Extents
{
    std::array<int, N> dims{}, strides{}, offsets{};
    int ncells = 0;
    Extents() = default;
    explicit Extents(const std::array<int, N>& d) { reset(d); }
    void reset(const std::array<int, N>& d)
    {
        dims = d; offsets = {};
        long n = 1;
        for (int k = 0; k < N; ++k) { strides[k] = (int)n; n *= d[k]; }
        ncells = (int)n;
    }
    void clear() { dims = {}; strides = {}; offsets = {}; ncells = 0; }
    int at(const std::array<int, N>& idx) const
    {
        int pos = 0;
        for (int k = 0; k < N; ++k) pos += (idx[k] - offsets[k] - 1) * strides[k];
        return pos;
    }
};

// what Array::subset / Array_gpu::subset extract: per dimension the first index and the extent;
// a source dimension of extent 1 is broadcast (include/Array.h:236-260, 579-622)
template<int N> struct Window
{
    std::array<int, N> first{}, extent{};
    explicit Window(const std::array<std::pair<int, int>, N>& ranges)
    {
        for (int k = 0; k < N; ++k) { first[k] = ranges[k].first; extent[k] = ranges[k].second - ranges[k].first + 1; }
    }
};
}  // namespace rrtmgp_b200

template<typename T, int N> class Array_gpu;

template<typename T, int N>
class Array
{
    public:
        Array() = default;
        Array(const std::array<int, N>& dims) : e(dims), data((size_t)e.ncells) {}                      // zeros
        Array(const std::vector<T>& values, const std::array<int, N>& dims)                               // copies at most ncells
            : e(dims), data(values.begin(), values.begin() + std::min<size_t>(values.size(), (size_t)e.ncells)) { data.resize((size_t)e.ncells); }
        Array(std::vector<T>&& values, const std::array<int, N>& dims) : e(dims), data(std::move(values)) {}
        Array(const Array&) = default;
        Array& operator=(const Array&) = default;
        Array(Array&& o) noexcept : e(o.e), data(std::move(o.data)) { o.e.clear(); o.data.clear(); }      // the source becomes empty
        Array& operator=(Array&& o) noexcept
        {
            if (this != &o) { e = o.e; data = std::move(o.data); o.e.clear(); o.data.clear(); }
            return *this;
        }
        Array(const Array_gpu<T, N>& device);                                                             // blocking D2H (Array.h:147-157)

        void set_offsets(const std::array<int, N>& offsets) { e.offsets = offsets; }
        std::array<int, N> get_dims() const { return e.dims; }
        void set_dims(const std::array<int, N>& dims)
        {
            if (e.ncells != 0) throw std::runtime_error("Only arrays of size 0 can be resized");
            e.reset(dims); data.resize((size_t)e.ncells);
        }
        void expand_dims(const std::array<int, N>& dims)
        {
            if (e.ncells == 0) throw std::runtime_error("Only arrays of size > 0 can be expanded");
            rrtmgp_b200::Extents<N> ne(dims);
            if (data.size() > (size_t)ne.ncells) throw std::runtime_error("new dimensions too small for data");
            e = ne; data.resize((size_t)e.ncells);
        }
        std::vector<T>& v() { return data; }
        const std::vector<T>& v() const { return data; }
        T* ptr() { return data.data(); }
        const T* ptr() const { return data.data(); }
        int size() const { return e.ncells; }
        T max() const { return *std::max_element(data.begin(), data.end()); }
        T min() const { return *std::min_element(data.begin(), data.end()); }
        void operator=(std::vector<T>&& values) { data = std::move(values); }
        T& operator()(const std::array<int, N>& indices) { return data[(size_t)e.at(indices)]; }
        T operator()(const std::array<int, N>& indices) const { return data[(size_t)e.at(indices)]; }
        int dim(const int i) const { return e.dims[i - 1]; }
        bool is_empty() const { return e.ncells == 0; }
        void fill(const T value) { std::fill(data.begin(), data.end(), value); }

        Array subset(const std::array<std::pair<int, int>, N> ranges) const
        {
            const rrtmgp_b200::Window<N> w(ranges);
            Array out(w.extent);
            std::array<int, N> idx;
            for (int pos = 0; pos < out.e.ncells; ++pos)
            {
                int rest = pos;
                for (int k = 0; k < N; ++k)                          // column-major: first index fastest
                {
                    const int ik = rest % w.extent[k];
                    rest /= w.extent[k];
                    idx[k] = e.dims[k] == 1 ? 1 : w.first[k] + ik;
                }
                out.data[(size_t)pos] = (*this)(idx);
            }
            return out;
        }

        void dump(const std::string& name) const
        {
            const std::string file_name = name + ".bin";
            std::ofstream f(file_name, std::ios::out | std::ios::trunc | std::ios::binary);
            if (!f) throw std::runtime_error("Cannot write file \"" + file_name + "\"");
            f.write(reinterpret_cast<const char*>(data.data()), (std::streamsize)(data.size() * sizeof(T)));
        }

    private:
        rrtmgp_b200::Extents<N> e;
        std::vector<T> data;
        friend class Array_gpu<T, N>;
};

// std::vector<bool> has no data(): Array<bool,N> is not supported, like Bool = signed char in the reference (types.h).

template<typename T, int N>
class Array_gpu
{
    public:
        Array_gpu() = default;
        Array_gpu(const std::array<int, N>& dims) : e(dims) { allocate(); }                               // uninitialised device memory
        Array_gpu(T* ptr, const std::array<int, N>& dims) : e(dims), p(ptr), view(true) {}                // view on caller memory
        Array_gpu(const Array<T, N>& host) : e(host.e) { allocate(); upload(host.ptr()); }                // H2D (Array.h:489-500)
        Array_gpu(const std::vector<T>& host, const std::array<int, N>& dims) : e(dims)                   // extension: from a flat vector
        {
            if (host.size() != (size_t)e.ncells) throw std::runtime_error("Array_gpu: size mismatch");
            allocate(); upload(host.data());
        }
        Array_gpu(const Array_gpu& o) : e(o.e)                                                            // deep copy; a view stays a view
        {
            if (o.view) { p = o.p; view = true; }
            else { allocate(); d2d(o.p); }
        }
        Array_gpu(Array_gpu&& o) noexcept : e(o.e), p(o.p), view(o.view) { o.e.clear(); o.p = nullptr; o.view = false; }
        ~Array_gpu() { release(); }

        Array_gpu& operator=(const Array_gpu& o)
        {
            if (this == &o) return *this;
            if (!(e.ncells == o.e.ncells || (e.ncells == 0 && p == nullptr)))
                throw std::runtime_error("Initialised arrays can not be resized");
            const bool fresh = e.ncells == 0;
            e = o.e;
            if (o.view) { release_keep_shape(); p = o.p; view = true; }
            else { if (fresh) allocate(); d2d(o.p); }
            return *this;
        }
        Array_gpu& operator=(Array_gpu&& o) noexcept(false)
        {
            if (this == &o) return *this;
            if (!(e.ncells == o.e.ncells || (e.ncells == 0 && p == nullptr)))
                throw std::runtime_error("initialised arrays can not be resized");
            release();
            e = o.e; p = o.p; view = o.view;
            o.e.clear(); o.p = nullptr; o.view = false;
            return *this;
        }

        void set_offsets(const std::array<int, N>& offsets) { e.offsets = offsets; }
        std::array<int, N> get_dims() const { return e.dims; }
        void set_dims(const std::array<int, N>& dims)
        {
            if (!(e.ncells == 0 && p == nullptr)) throw std::runtime_error("Only arrays with uninitialized pointers can be resized");
            e.reset(dims); allocate();
        }
        void set_data(const Array<T, N>& host) { if (!p) allocate(); upload(host.ptr()); }
        void fill(const T value)
        {
            if (e.ncells) rrtmgp_b200::check(rrtmgp_b200_array_fill((int)sizeof(T), (long long)e.ncells, p, &value, rrtmgp_b200::stream()));
        }
        // single elements (Array.h:543-573): device-to-device, host-to-device, device-to-host
        void copy(const std::array<int, N>& indices, Array_gpu& input, const std::array<int, N>& indices_input) const
        {
            rrtmgp_b200::cuda_check(cudaMemcpyAsync(p + e.at(indices), input.p + input.e.at(indices_input), sizeof(T),
                                                    cudaMemcpyDeviceToDevice, rrtmgp_b200::stream()));
        }
        void insert(const std::array<int, N>& indices, const T value) const
        {
            rrtmgp_b200::cuda_check(cudaMemcpyAsync(p + e.at(indices), &value, sizeof(T), cudaMemcpyHostToDevice, rrtmgp_b200::stream()));
            rrtmgp_b200::cuda_check(cudaStreamSynchronize(rrtmgp_b200::stream()));      // `value` is a stack temporary
        }
        T operator()(const std::array<int, N>& indices) const
        {
            T value;
            rrtmgp_b200::cuda_check(cudaMemcpyAsync(&value, p + e.at(indices), sizeof(T), cudaMemcpyDeviceToHost, rrtmgp_b200::stream()));
            rrtmgp_b200::cuda_check(cudaStreamSynchronize(rrtmgp_b200::stream()));
            return value;
        }
        T* ptr() { return p; }
        const T* ptr() const { return p; }
        int size() const { return e.ncells; }
        int dim(const int i) const { return e.dims[i - 1]; }

        // a new owning array holding the window; source dimensions of extent 1 are broadcast.  One kernel on the
        // stream, no synchronisation (the reference: include/Array.h:579-622, 64-thread blocks on the default stream).
        Array_gpu subset(const std::array<std::pair<int, int>, N> ranges) const
        {
            const rrtmgp_b200::Window<N> w(ranges);
            Array_gpu out(w.extent);
            window_into(out, w.first);
            return out;
        }
        // fills the caller's array in place; the return value is a view of it (the reference returns a deep copy)
        Array_gpu subset_copy(Array_gpu& a_sub, const std::array<int, N>& block_corners) const
        {
            window_into(a_sub, block_corners);
            return Array_gpu(a_sub.p, a_sub.e.dims);
        }
        void dump(const std::string& name) const { Array<T, N>(*this).dump(name); }

        // extensions used by this repository's drivers and tests
        void copy_from(const std::vector<T>& h)
        {
            if (h.size() != (size_t)e.ncells) throw std::runtime_error("Array_gpu: size mismatch");
            upload(h.data());
        }
        std::vector<T> to_host() const
        {
            std::vector<T> h((size_t)e.ncells);
            if (e.ncells)
            {
                rrtmgp_b200::cuda_check(cudaMemcpyAsync(h.data(), p, h.size() * sizeof(T), cudaMemcpyDeviceToHost, rrtmgp_b200::stream()));
                rrtmgp_b200::cuda_check(cudaStreamSynchronize(rrtmgp_b200::stream()));
            }
            return h;
        }

    private:
        void window_into(Array_gpu& a_sub, const std::array<int, N>& first) const
        {
            if (a_sub.e.ncells)
                rrtmgp_b200::check(rrtmgp_b200_array_subset((int)sizeof(T), N, e.dims.data(), e.offsets.data(), first.data(),
                                                            a_sub.e.dims.data(), p, a_sub.p, rrtmgp_b200::stream()));
        }
        void allocate()
        {
            view = false; p = nullptr;
            if (e.ncells) rrtmgp_b200::cuda_check(cudaMallocAsync((void**)&p, (size_t)e.ncells * sizeof(T), rrtmgp_b200::stream()));
        }
        void upload(const T* h)          // pageable host memory: the copy is staged before the call returns
        {
            if (e.ncells) rrtmgp_b200::cuda_check(cudaMemcpyAsync(p, h, (size_t)e.ncells * sizeof(T), cudaMemcpyHostToDevice, rrtmgp_b200::stream()));
        }
        void d2d(const T* src)
        {
            if (e.ncells) rrtmgp_b200::cuda_check(cudaMemcpyAsync(p, src, (size_t)e.ncells * sizeof(T), cudaMemcpyDeviceToDevice, rrtmgp_b200::stream()));
        }
        void release_keep_shape() { if (!view && p) cudaFreeAsync(p, rrtmgp_b200::stream()); p = nullptr; view = false; }
        void release() { release_keep_shape(); e.clear(); }
        rrtmgp_b200::Extents<N> e;
        T* p = nullptr;
        bool view = false;
        friend class Array<T, N>;
};

template<typename T, int N>
Array<T, N>::Array(const Array_gpu<T, N>& device) : e(device.e), data((size_t)device.e.ncells)
{
    if (e.ncells)
    {
        rrtmgp_b200::cuda_check(cudaMemcpyAsync(data.data(), device.ptr(), data.size() * sizeof(T), cudaMemcpyDeviceToHost, rrtmgp_b200::stream()));
        rrtmgp_b200::cuda_check(cudaStreamSynchronize(rrtmgp_b200::stream()));
    }
}

// include/Array.h:652-670
template<typename T, int N>
bool any_vals_outside(const Array<T, N>& array, const T lower_limit, const T upper_limit)
{
    for (const T val : array.v()) if (val < lower_limit || val > upper_limit) return true;
    return false;
}
template<typename T, int N>
bool any_vals_less_than(const Array<T, N>& array, const T lower_limit)
{
    for (const T val : array.v()) if (val < lower_limit) return true;
    return false;
}
