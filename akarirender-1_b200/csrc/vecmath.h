// vecmath.h — fp32 vector helpers shared by host (scene flattening, BVH build) and device code.
//
// Parity contract: IEEE fp32, NO fused multiply-add contraction, every expression in
// the reference's operator order. Device code is compiled with `-fmad=false`, host code with
// `-ffp-contract=off`; akb_init() runs a self-test kernel that fails loudly if contraction is on.
// Each helper cites the reference function whose arithmetic it restates (paths under
// /root/reference/src/akari/).
#pragma once
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define AKB_HD __host__ __device__ __forceinline__
#else
#define AKB_HD inline
#endif

namespace akb {

struct v3 {
    float x, y, z;
};

AKB_HD v3 mk3(float x, float y, float z) {
    v3 r;
    r.x = x, r.y = y, r.z = z;
    return r;
}
// common/array.h:203-210 — dot: s = a0*b0; s += a1*b1; s += a2*b2
AKB_HD float dot3(v3 a, v3 b) {
    float s = a.x * b.x;
    s += a.y * b.y;
    s += a.z * b.z;
    return s;
}
// common/math.h:177-181
AKB_HD v3 cross3(v3 a, v3 b) { return mk3((a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x)); }
AKB_HD v3 add3(v3 a, v3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
AKB_HD v3 sub3(v3 a, v3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
AKB_HD v3 mul3(v3 a, v3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
AKB_HD v3 scale3(float s, v3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
AKB_HD v3 div3s(v3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
AKB_HD v3 neg3(v3 a) { return mk3(-a.x, -a.y, -a.z); }
// common/array.h:283-286 — normalize: a / sqrt(dot(a,a)), three true divisions
AKB_HD v3 normalize3(v3 a) { return div3s(a, sqrtf(dot3(a, a))); }
// std::max / std::min semantics (common/array.h:41-46): max(a,b) = (a<b)?b:a ; min(a,b) = (b<a)?b:a
AKB_HD float max_std(float a, float b) { return (a < b) ? b : a; }
AKB_HD float min_std(float a, float b) { return (b < a) ? b : a; }
// common/math.h:49-51 — lerp3: (1-u-v)*a + u*b + v*c
AKB_HD v3 lerp3(v3 a, v3 b, v3 c, float u, float v) {
    float w = 1.0f - u - v;
    return add3(add3(scale3(w, a), scale3(u, b)), scale3(v, c));
}

// common/math.h:34-42
#define AKB_PI ((float)3.1415926535897932384f)
#define AKB_PI2 (AKB_PI / 2.0f)
#define AKB_PI4 (AKB_PI / 4.0f)
#define AKB_INVPI (1.0f / AKB_PI)
#define AKB_EPS 0.001f
#define AKB_SHADOW_EPS 0.0001f

} // namespace akb
