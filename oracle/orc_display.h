// orc_display.h — CPU restatement of the reference's presentation pass (TEST INFRASTRUCTURE ONLY, like the rest of oracle/).
//
// Follows renderer/Shaders.metal:33-69 (blitFragment: nearest sample of the accumulated image, optional tone mapping
// 1 - exp(-c), optional manual sRGB encode, then one of the four comparison modes against a reference image) and
// renderer/Raytracing.h:130-135 (toSRGB), with the switches of renderer/Raytracing.h:11-12, 27-33 turned from macros
// into parameters.  The full-screen triangle of Shaders.metal:18-31 maps pixel centre (x + .5, y + .5) / size to the
// sample coordinate, so an output of the image's own size reads pixel (x, y) of the image and the nearest texel of a
// reference of any size.  Parity status: exp/pow come from libm here and from libdevice on the GPU — compared within a
// stated tolerance (tests/test_display.py), everything else bit-exact.
#pragma once
#include <cmath>
#include <cstdint>

namespace orc {

enum : uint32_t { COMPARE_DISABLED = 0, COMPARE_ABSOLUTE_VALUE = 1, COMPARE_REF_TO_COLOR = 2, COMPARE_COLOR_TO_REF = 3, COMPARE_LUMINANCE = 4 };

struct DisplayParams { uint32_t toneMapping, manualSRGB, comparisonMode; float comparisonScale; };

static inline float toSRGB(float v) {                               // Raytracing.h:130-135
    if (v <= 0.0f) return 0.0f;
    if (v >= 1.0f) return 1.0f;
    return (v < 0.0031308f) ? (12.92f * v) : (1.055f * std::pow(v, 1.0f / 2.4f) - 0.055f);
}

static inline uint32_t nearest_texel(uint32_t i, uint32_t n, uint32_t m) {     // texel of an m-wide texture under pixel i of n
    float u = (float(i) + 0.5f) / float(n);
    uint32_t t = uint32_t(u * float(m));
    return t < m ? t : m - 1;
}

// image: W*H RGBA (row 0 = bottom, as accumulated); ref: RW*RH RGBA or null; out: W*H RGBA
static inline void display(const float* image, uint32_t W, uint32_t H, const DisplayParams& p, const float* ref, uint32_t RW,
                           uint32_t RH, float* out) {
    for (uint32_t y = 0; y < H; ++y)
        for (uint32_t x = 0; x < W; ++x) {
            float c[4], r[4] = {0, 0, 0, 0};
            for (int k = 0; k < 4; ++k) c[k] = image[4 * (size_t(y) * W + x) + k];
            if (p.toneMapping) for (int k = 0; k < 4; ++k) c[k] = 1.0f - std::exp(-c[k]);          // Shaders.metal:43-45
            if (p.manualSRGB) for (int k = 0; k < 3; ++k) c[k] = toSRGB(c[k]);                     // Shaders.metal:47-51
            if (p.comparisonMode != COMPARE_DISABLED && ref) {
                size_t at = size_t(nearest_texel(y, H, RH)) * RW + nearest_texel(x, W, RW);
                for (int k = 0; k < 4; ++k) r[k] = ref[4 * at + k];
            }
            float* o = out + 4 * (size_t(y) * W + x);
            const float s = p.comparisonScale;
            switch (ref ? p.comparisonMode : COMPARE_DISABLED) {
            case COMPARE_ABSOLUTE_VALUE: for (int k = 0; k < 4; ++k) o[k] = std::fabs(c[k] - r[k]) * s; break;          // :53-55
            case COMPARE_REF_TO_COLOR:   for (int k = 0; k < 4; ++k) o[k] = std::fmax(0.0f, r[k] - c[k]) * s; break;    // :56-58
            case COMPARE_COLOR_TO_REF:   for (int k = 0; k < 4; ++k) o[k] = std::fmax(0.0f, c[k] - r[k]) * s; break;    // :59-61
            case COMPARE_LUMINANCE: {                                                                                   // :62-66
                const float third = 1.0f / 3.0f;
                float lc = std::fmaf(c[2], third, std::fmaf(c[1], third, c[0] * third));
                float lr = std::fmaf(r[2], third, std::fmaf(r[1], third, r[0] * third));
                o[0] = std::fmax(0.0f, lc - lr) * s; o[1] = std::fmax(0.0f, lr - lc) * s; o[2] = 0.0f * s; o[3] = 1.0f * s;
                break;
            }
            default: for (int k = 0; k < 4; ++k) o[k] = c[k]; break;                                                   // :67-68
            }
        }
}

}  // namespace orc
