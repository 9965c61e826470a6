// Minimal stand-in for Apple's <simd/simd.h>, enough for the HOST half of the reference's
// renderer/Raytracing.h + Spectrum.h to compile with g++ (used by ref_probe.cpp only).
#pragma once
typedef float vector_float4 __attribute__((vector_size(16)));   // 16-byte aligned, like simd
typedef float vector_float3 __attribute__((vector_size(16)));
struct packed_float2 { float x, y; };
