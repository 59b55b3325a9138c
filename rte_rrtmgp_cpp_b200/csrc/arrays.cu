// arrays.cu — the two kernels behind Array_gpu::fill and Array_gpu::subset (include/rrtmgp_b200/array.hpp; reference
// include/Array.h:311-350 fill_kernel / subset_kernel, 64-thread blocks on the default stream).  Type-erased by element
// size (1, 2, 4, 8, 16 bytes), stream-ordered, grid-stride with a grid sized to the device.
#include "stages.hpp"
#include "launch.hpp"

namespace rb200 {

template<int BYTES> struct Word;
template<> struct Word<1> { using type = unsigned char; };
template<> struct Word<2> { using type = unsigned short; };
template<> struct Word<4> { using type = unsigned int; };
template<> struct Word<8> { using type = unsigned long long; };
template<> struct Word<16> { using type = uint4; };

template<typename W> __global__ void array_fill_kernel(W* __restrict__ a, const W v, const size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) a[i] = v;
}

// out(j1..jN) = in(first + j), a source dimension of extent 1 broadcast; both column-major.  The output index is
// decomposed fastest dimension first, so consecutive threads read consecutive source elements along dimension 1.
template<typename W> __global__ void array_subset_kernel(W* __restrict__ out, const W* __restrict__ in, const SubsetGeom g, const size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    {
        size_t rest = i, src = 0;
#pragma unroll
        for (int k = 0; k < SubsetGeom::MAXD; ++k)
        {
            if (k < g.ndim)
            {
                const size_t jk = rest % (size_t)g.extent[k];
                rest /= (size_t)g.extent[k];
                src += (g.spread[k] ? (size_t)0 : jk + (size_t)g.start0[k]) * g.stride[k];
            }
        }
        out[i] = in[src];
    }
}

static unsigned grid_for(size_t n)
{
    const size_t blocks = (n + 255) / 256;
    return (unsigned)(blocks < 1 ? 1 : (blocks > 148 * 16 ? 148 * 16 : blocks));
}

template<int BYTES> static void fill_t(void* a, const void* value, size_t n, cudaStream_t s)
{
    using W = typename Word<BYTES>::type;
    W v;
    memcpy(&v, value, BYTES);
    array_fill_kernel<W><<<grid_for(n), 256, 0, s>>>((W*)a, v, n);
    count_launch();
}
template<int BYTES> static void subset_t(void* out, const void* in, const SubsetGeom& g, size_t n, cudaStream_t s)
{
    using W = typename Word<BYTES>::type;
    array_subset_kernel<W><<<grid_for(n), 256, 0, s>>>((W*)out, (const W*)in, g, n);
    count_launch();
}

bool launch_array_fill(int elem_bytes, void* a, const void* value, size_t n, cudaStream_t s)
{
    switch (elem_bytes)
    {
    case 1: fill_t<1>(a, value, n, s); return true;
    case 2: fill_t<2>(a, value, n, s); return true;
    case 4: fill_t<4>(a, value, n, s); return true;
    case 8: fill_t<8>(a, value, n, s); return true;
    case 16: fill_t<16>(a, value, n, s); return true;
    default: return false;
    }
}
bool launch_array_subset(int elem_bytes, void* out, const void* in, const SubsetGeom& g, size_t n, cudaStream_t s)
{
    switch (elem_bytes)
    {
    case 1: subset_t<1>(out, in, g, n, s); return true;
    case 2: subset_t<2>(out, in, g, n, s); return true;
    case 4: subset_t<4>(out, in, g, n, s); return true;
    case 8: subset_t<8>(out, in, g, n, s); return true;
    case 16: subset_t<16>(out, in, g, n, s); return true;
    default: return false;
    }
}

}  // namespace rb200
