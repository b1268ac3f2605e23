// The 16-bit operand format of the tensor-core path.
//
// Every activation / weight tensor the tcgen05 kernels read is 16 bits per element with fp32 accumulation in TMEM
// (`tcgen05.mma.kind::f16` takes fp16 or bf16 operands at the same rate; the instruction descriptor selects which).
// A handle uses ONE format for all of its tensors:
//   fp16 (default): 11 significant bits.  D(G(z)) stays within 2e-3 of the float64 oracle on every architecture
//                   (BASELINE north_star asks for 1e-2), x_mean within 5e-3, at no cost in throughput.
//   bf16          : 8 significant bits, fp32's exponent range (apvae_config.operand_bf16 = 1) for checkpoints whose
//                   activations could leave fp16's range (|v| < 65504); energies then sit at ~1e-2 (tools/explore_parity.py).
// Overflow in fp16 is not clamped: it propagates as inf / nan into the metrics (loud, not silent).
#pragma once
#include <stdint.h>
#include <string.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

namespace apvae {
namespace op16 {

// ---- host (weight packing; same rounding as the device conversions: round to nearest even)
inline uint16_t host_from_f32(float f, bool f16) {
    if (f16) {
        return static_cast<__half_raw>(__float2half_rn(f)).x;
    }
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7F800000u) == 0x7F800000u && (u & 0x007FFFFFu)) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7FFFu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}
inline float host_to_f32(uint16_t v, bool f16) {
    if (f16) {
        return __half2float(__half(__half_raw{v}));
    }
    const uint32_t u = (uint32_t)v << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}
inline float host_round(float f, bool f16) { return host_to_f32(host_from_f32(f, f16), f16); }

#ifdef __CUDACC__
// ---- device: two values per 32-bit word (element 0 in the low half)
template <bool F16>
__device__ __forceinline__ uint32_t pack(float a, float b) {
    if constexpr (F16) {
        const __half2 h = __floats2half2_rn(a, b);
        return *reinterpret_cast<const uint32_t*>(&h);
    } else {
        const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<const uint32_t*>(&h);
    }
}
template <bool F16>
__device__ __forceinline__ float lo(uint32_t u) {
    if constexpr (F16) return __half2float(__ushort_as_half((unsigned short)(u & 0xffffu)));
    else return __uint_as_float(u << 16);
}
template <bool F16>
__device__ __forceinline__ float hi(uint32_t u) {
    if constexpr (F16) return __half2float(__ushort_as_half((unsigned short)(u >> 16)));
    else return __uint_as_float(u & 0xffff0000u);
}
template <bool F16>
__device__ __forceinline__ float2 unpack(uint32_t u) {
    if constexpr (F16) return __half22float2(*reinterpret_cast<const __half2*>(&u));
    else return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
}
template <bool F16>
__device__ __forceinline__ float round(float a) {
    if constexpr (F16) return __half2float(__float2half_rn(a));
    else return __bfloat162float(__float2bfloat16_rn(a));
}
template <bool F16>
__device__ __forceinline__ uint16_t from_f32(float a) {
    if constexpr (F16) return __half_as_ushort(__float2half_rn(a));
    else return __bfloat16_as_ushort(__float2bfloat16_rn(a));
}
// run-time format (HBM-bound CUDA-core kernels)
__device__ __forceinline__ uint16_t from_f32(float a, bool f16) { return f16 ? from_f32<true>(a) : from_f32<false>(a); }
__device__ __forceinline__ float round(float a, bool f16) { return f16 ? round<true>(a) : round<false>(a); }
// "> 0" of a stored element, whatever the format: sign bit clear and not zero (leaky-relu' masks)
__device__ __forceinline__ bool pos_lo(uint32_t u) { return (int16_t)(u & 0xffffu) > 0; }
__device__ __forceinline__ bool pos_hi(uint32_t u) { return (int32_t)u > 0xffff; }
#endif

}  // namespace op16
}  // namespace apvae
