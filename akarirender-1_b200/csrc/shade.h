// shade.h — per-vertex math of the unidirectional path tracer: sampler, camera, BSDF closures, area-light NEE.
//
// Restates, operation for operation, what GenericPathTracer<C>::run_megakernel does at one path vertex
// (kernel/pathtracer.h:61-163) and everything it calls. Each function cites the reference lines it follows.
// fp32, no FMA contraction (compile flags, vecmath.h), libm calls evaluated so that they round like the host's
// (AKB_EXACT_LIBM). AKB_HD: the same code runs in the CUDA kernels and in tests/hostcheck.
#pragma once
#include "../../include/akari_b200.h"
#include "traverse.h"

namespace akb {

#ifndef AKB_EXACT_LIBM
#define AKB_EXACT_LIBM 1
#endif

// The reference's host build calls glibc's sinf/cosf. Those are NOT correctly rounded (about 1.3 % of arguments
// differ from the correctly rounded value by one ulp), so bit parity with the reference's images needs the same
// algorithm, not a more accurate one. glibc >= 2.28 (sysdeps/ieee754/flt-32/s_sinf.c, s_cosf.c, sincosf.h — the ARM
// optimized-routines sincosf; a third-party dependency of the reference, not part of /root/reference) evaluates in
// double: |y| < ~pi/4 directly, |y| < 120 after the 2/pi * 2^24 fixed-point quadrant reduction, with one degree-7 /
// degree-8 polynomial pair. x86-64 glibc selects its FMA build at run time (ifunc), where every a + b*c of that source
// is a fused multiply-add; the restatement below uses fma() in exactly those places. Checked against glibc 2.39 on
// 10^8 arguments: 0 mismatches (without the FMAs: 3). Arguments here never reach 120 (theta in [-pi/4, 3pi/4], phi in
// [0, 2pi]), so the large-argument branch is not needed; it falls back to the double-precision routine.
struct SinCosTab {
    double c0, c1, c2, c3, c4, s1, s2, s3;
};
AKB_HD uint32_t abstop12(float x) { return (akb_f2u(x) >> 20) & 0x7ffu; }
AKB_HD float sincosf_poly(double x, double x2, bool neg, int n) {
    // table entry 1 (used when the quadrant has bit 1 set) is entry 0 with the cosine coefficients negated
    const double sg = neg ? -1.0 : 1.0;
    const SinCosTab p = {sg * 0x1p0, sg * -0x1.ffffffd0c621cp-2, sg * 0x1.55553e1068f19p-5, sg * -0x1.6c087e89a359dp-10,
                         sg * 0x1.99343027bf8c3p-16, -0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13};
    if ((n & 1) == 0) {
        double x3 = x * x2;
        double s1 = fma(x2, p.s3, p.s2);
        double x7 = x3 * x2;
        double s = fma(x3, p.s1, x);
        return (float)fma(x7, s1, s);
    } else {
        double x4 = x2 * x2;
        double c2 = fma(x2, p.c4, p.c3);
        double c1 = fma(x2, p.c1, p.c0);
        double x6 = x4 * x2;
        double c = fma(x4, p.c2, c1);
        return (float)fma(x6, c2, c);
    }
}
AKB_HD float sincosf_glibc(float y, int is_cos) {
    double x = (double)y;
    if (abstop12(y) < 0x3f4u) { // abstop12(pio4f)
        if (abstop12(y) < 0x398u) return is_cos ? 1.0f : y; // |y| < 2^-12
        return sincosf_poly(x, x * x, false, is_cos);
    }
    if (abstop12(y) < 0x42fu) { // abstop12(120.0f)
        const double r = x * 0x1.45F306DC9C883p+23;
        const int n = ((int32_t)r + 0x800000) >> 24;
        x = fma(-(double)n, 0x1.921FB54442D18p0, x);
        const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0; // sign[] = {1,-1,-1,1}
        return sincosf_poly(x * sgn, x * x, (n & 2) != 0, n ^ is_cos);
    }
    return is_cos ? (float)cos(x) : (float)sin(x);
}
// sinf(y) and cosf(y) of the same argument: the two calls above with their common part - the quadrant reduction -
// written once. Each result is the value sincosf_glibc returns for it, statement for statement.
AKB_HD void sincosf_glibc_both(float y, float &s, float &c) {
    double x = (double)y;
    if (abstop12(y) < 0x3f4u) {
        if (abstop12(y) < 0x398u) {
            s = y, c = 1.0f;
            return;
        }
        const double x2 = x * x;
        s = sincosf_poly(x, x2, false, 0), c = sincosf_poly(x, x2, false, 1);
        return;
    }
    if (abstop12(y) < 0x42fu) {
        const double r = x * 0x1.45F306DC9C883p+23;
        const int n = ((int32_t)r + 0x800000) >> 24;
        x = fma(-(double)n, 0x1.921FB54442D18p0, x);
        const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
        const double xs = x * sgn, x2 = x * x;
        const bool neg = (n & 2) != 0;
        s = sincosf_poly(xs, x2, neg, n), c = sincosf_poly(xs, x2, neg, n ^ 1);
        return;
    }
    s = (float)sin(x), c = (float)cos(x);
}
#if defined(__CUDA_ARCH__) && AKB_EXACT_LIBM
AKB_HD float sin_ref(float x) { return sincosf_glibc(x, 0); }
AKB_HD float cos_ref(float x) { return sincosf_glibc(x, 1); }
AKB_HD void sincos_ref(float x, float &s, float &c) { sincosf_glibc_both(x, s, c); }
#else
AKB_HD float sin_ref(float x) { return sinf(x); }
AKB_HD float cos_ref(float x) { return cosf(x); }
AKB_HD void sincos_ref(float x, float &s, float &c) { s = sinf(x), c = cosf(x); }
#endif

// ---- flattened scene as the kernels see it (all pointers into HBM) ----------------------------------------
struct CameraDev {
    float c2w[16], r2c[16];
    float ox, oy, oz; // c2w.apply_point(0,0,0), computed once on the host with apply_point's arithmetic
    int width, height;
};

struct SceneDev {
    const float4 *nodes;   // 5 x float4 per Node8
    const float4 *tris;    // 3 x float4 per triangle, leaf order: (v0|flat id) (e1|material) (e2|0)
    const float4 *shading; // 6 x float4 per triangle, flat order:
                           // (v0|material) (v1|ng.x) (v2|ng.y) (n0|ng.z) (n1|geom id) (n2|prim id)
    const akb_material *mats;
    const float4 *lights;  // 4 x float4 per area light: (v0|area) (v1|Le.r) (v2|Le.g) (ng|Le.b)
    const float *light_cdf;  // n_lights + 1
    const float *light_func; // n_lights
    float light_func_int;
    int n_lights;
    int n_mats;
    int trav_post_limit; // traversal stack height below which a leftover triangle group may be parked (trav_step_ii)
    uint32_t n_nodes, n_tris; // sizes of nodes / tris (bounds checks of the -DAKB_DEBUG_CHECKS build)
    // image textures (kernel/texture.h:39-57); n_textures == 0 for every BASELINE config, and then none of this is read
    int n_textures;
    const float4 *texels;   // all textures back to back, RGBA
    const int4 *tex_info;   // (first texel, width, height, -)
    const float4 *tri_tc;   // 2 x float4 per triangle, flat order: (u0,v0,u1,v1) (u2,v2,-,-)
    const float4 *light_tc; // 2 x float4 per light: (u0,v0,u1,v1) (u2,v2,texture id bits,-)
    CameraDev cam;
};

// ---- samplers: kernel/sampler.h:28-66 -------------------------------------------------------------------------
// SAMP = AKB_SAMPLER_LCG: state in the low 32 bits; AKB_SAMPLER_PCG: 64-bit state.
template <int SAMP> struct Rng {
    uint64_t s;
    AKB_HD uint32_t pcg32() { // :33-41
        uint64_t x = s;
        unsigned count = (unsigned)(x >> 59);
        s = x * 6364136223846793005ull + 1442695040888963407ull;
        x ^= x >> 18;
        uint32_t v = (uint32_t)(x >> 27);
        return v >> count | v << (-count & 31);
    }
    AKB_HD void set_sample_index(uint64_t idx) {
        if (SAMP == AKB_SAMPLER_PCG) { // :42-45
            s = idx + 1442695040888963407ull;
            (void)pcg32();
        } else { // :59
            s = idx & 0xffffffffull;
        }
    }
    AKB_HD float next1d() {
        if (SAMP == AKB_SAMPLER_PCG) return (float)pcg32() / (float)0xffffffffu; // :50 ((float)0xffffffff == 2^32)
        uint32_t seed = (uint32_t)s;
        seed = 1103515245u * seed + 12345u; // :60-63
        s = seed;
        return (float)seed / (float)0xFFFFFFFFu;
    }
    // next2d() = float2(next1d(), next1d()) (:51,:64): the oracle build (g++ 13) evaluates right to left, so y gets
    // the FIRST draw and x the SECOND (pinned by the KAT in tests/golden).
    AKB_HD void next2d(float &x, float &y) {
        y = next1d();
        x = next1d();
    }
};

// ---- camera: kernel/camera.h:67-86, Transform::apply_point common/math.h:251-259 --------------------------
AKB_HD v3 apply_point(const float *m, v3 p) {
    float v[4];
    for (int i = 0; i < 4; i++) {
        float s = m[4 * i + 0] * p.x;
        s += m[4 * i + 1] * p.y;
        s += m[4 * i + 2] * p.z;
        s += m[4 * i + 3] * 1.0f;
        v[i] = s;
    }
    v3 q = mk3(v[0], v[1], v[2]);
    if (v[3] != 1.0f) q = div3s(q, v[3]);
    return q;
}
AKB_HD void camera_generate_ray(const CameraDev &c, float u2x, float u2y, int px, int py, v3 &o, v3 &d) {
    float fx = (float)px + u2x, fy = (float)py + u2y;
    v3 p = apply_point(c.r2c, mk3(fx, fy, 0.0f));
    v3 dl = normalize3(sub3(mk3(p.x, p.y, 0), mk3(0, 0, 1)));
    o = mk3(c.ox, c.oy, c.oz);
    d.x = dot3(mk3(c.c2w[0], c.c2w[1], c.c2w[2]), dl); // apply_vector: m3 * v (common/math.h:261)
    d.y = dot3(mk3(c.c2w[4], c.c2w[5], c.c2w[6]), dl);
    d.z = dot3(mk3(c.c2w[8], c.c2w[9], c.c2w[10]), dl);
}

// ---- BSDF helpers: kernel/bsdf-funcs.h:26-89 (local frame is y-up) ----------------------------------------
AKB_HD bool same_hemisphere(v3 wo, v3 wi) { return wo.y * wi.y >= 0; }
AKB_HD float cos2_theta(v3 w) { return w.y * w.y; }
AKB_HD float sin2_theta(v3 w) { return 1 - cos2_theta(w); }
AKB_HD float tan2_theta(v3 w) { return sin2_theta(w) / cos2_theta(w); }
AKB_HD v3 reflect3(v3 w, v3 n) { return add3(scale3(-1.0f, w), scale3(2.0f * dot3(w, n), n)); } // :58-60
AKB_HD bool refract3(v3 wi, v3 n, float eta, v3 &wt) { // :62-73
    float cosThetaI = dot3(n, wi);
    float sin2ThetaI = fmaxf(0.0f, 1.0f - cosThetaI * cosThetaI);
    float sin2ThetaT = eta * eta * sin2ThetaI;
    if (sin2ThetaT >= 1) return false;
    float cosThetaT = sqrtf(1 - sin2ThetaT);
    wt = add3(scale3(eta, neg3(wi)), scale3(eta * cosThetaI - cosThetaT, n));
    return true;
}
AKB_HD float fr_dielectric(float cosThetaI, float etaI, float etaT) { // :74-89
    bool entering = cosThetaI > 0.f;
    if (!entering) {
        float x = etaI;
        etaI = etaT, etaT = x;
        cosThetaI = fabsf(cosThetaI);
    }
    float sinThetaI = sqrtf(max_std(0.0f, 1 - cosThetaI * cosThetaI));
    float sinThetaT = etaI / etaT * sinThetaI;
    if (sinThetaT >= 1) return 1;
    float cosThetaT = sqrtf(max_std(0.0f, 1 - sinThetaT * sinThetaT));
    float Rpar = ((etaT * cosThetaI) - (etaI * cosThetaT)) / ((etaT * cosThetaI) + (etaI * cosThetaT));
    float Rper = ((etaI * cosThetaI) - (etaT * cosThetaT)) / ((etaI * cosThetaI) + (etaT * cosThetaT));
    return (float)(0.5 * (double)(Rpar * Rpar + Rper * Rper));
}

// ---- sampling: kernel/sampling.h:32-69 ------------------------------------------------------------------------
AKB_HD void concentric_disk(float ux, float uy, float &ox, float &oy) {
    float x = 2.f * ux - 1, y = 2.f * uy - 1;
    if (x == 0 && y == 0) {
        ox = 0, oy = 0;
        return;
    }
    float theta, r;
    if (fabsf(x) > fabsf(y)) {
        r = x;
        theta = AKB_PI4 * (y / x);
    } else {
        r = y;
        theta = AKB_PI2 - AKB_PI4 * (x / y);
    }
    float st, ct;
    sincos_ref(theta, st, ct);
    ox = r * ct, oy = r * st;
}
AKB_HD v3 cosine_hemisphere(float ux, float uy) {
    float x, y;
    concentric_disk(ux, uy, x, y);
    float r = x * x;
    r += y * y;
    float h = sqrtf(max_std(0.0f, 1.0f - r));
    return mk3(x, h, y);
}

// ---- Frame: common/math.h:201-225 ----------------------------------------------------------------------------
struct Frame {
    v3 n, T, B;
};
AKB_HD Frame frame_make(v3 v1) {
    Frame f;
    f.n = v1;
    if (fabsf(v1.x) > fabsf(v1.y))
        f.T = div3s(mk3(-v1.z, 0, v1.x), sqrtf(v1.x * v1.x + v1.z * v1.z));
    else
        f.T = div3s(mk3(0, v1.z, -v1.y), sqrtf(v1.y * v1.y + v1.z * v1.z));
    f.B = normalize3(cross3(v1, f.T));
    return f;
}
AKB_HD v3 w2l(const Frame &f, v3 v) { return mk3(dot3(f.T, v), dot3(f.n, v), dot3(f.B, v)); }
AKB_HD v3 l2w(const Frame &f, v3 v) { return add3(add3(scale3(v.x, f.T), scale3(v.y, f.n)), scale3(v.z, f.B)); }

// ---- closures: kernel/material.h:57-138, kernel/microfacet.h:74-157 -------------------------------------------
enum { CL_DIFFUSE = 0, CL_GGX = 1, CL_DIELECTRIC = 2, CL_NULL = 3 };
struct Bsdf {
    int kind;
    v3 R;
    float alpha, ior, choice_pdf;
    Frame frame;
};

AKB_HD float ggx_D(float alpha, v3 m) { // microfacet.h:74-82
    if (m.y <= 0.0f) return 0.0f;
    float a2 = alpha * alpha, c2 = cos2_theta(m), t2 = tan2_theta(m), at = (a2 + t2);
    return a2 / (AKB_PI * c2 * c2 * at * at);
}
// microfacet.h:84-89: G1(v, m) = 0 if v and m disagree, else a term in the angle of m only, evaluated in double. The
// reference evaluates that term once per direction; it is the same number both times (same alpha, same m), so it is
// worked out once here (ggx_G1_term) and gated per direction - the double-precision square root and division are the
// largest block of the shade kernel.
AKB_HD float ggx_G1_term(float alpha, v3 m) { return (float)(2.0 / (1.0 + sqrt(1.0 + (double)(alpha * alpha * tan2_theta(m))))); }
AKB_HD float ggx_G1(float term, v3 v, v3 m) { return dot3(v, m) * v.y <= 0 ? 0.0f : term; }
AKB_HD v3 closure_evaluate(const Bsdf &b, v3 wo, v3 wi) {
    const v3 zero = mk3(0, 0, 0);
    if (b.kind == CL_DIFFUSE) { // material.h:69-74
        if (same_hemisphere(wo, wi)) return scale3(AKB_INVPI, b.R);
        return zero;
    }
    if (b.kind == CL_GGX) { // material.h:103-121
        if (!same_hemisphere(wo, wi)) return zero;
        float cosThetaO = fabsf(wo.y), cosThetaI = fabsf(wi.y);
        v3 wh = add3(wo, wi);
        if (cosThetaI == 0 || cosThetaO == 0) return zero;
        if (wh.x == 0 && wh.y == 0 && wh.z == 0) return zero;
        wh = normalize3(wh);
        if (wh.y < 0) wh = neg3(wh);
        float F = 1.0f;
        const float g1 = ggx_G1_term(b.alpha, wh);
        float s = ggx_D(b.alpha, wh) * (ggx_G1(g1, wo, wh) * ggx_G1(g1, wi, wh)) * F / (4.0f * cosThetaI * cosThetaO);
        return mk3(b.R.x * s, b.R.y * s, b.R.z * s);
    }
    return zero; // specular dielectric (extension) and the null closure evaluate to 0
}
// returns f; writes wi (local) and pdf
AKB_HD v3 closure_sample(const Bsdf &b, float ux, float uy, v3 wo, v3 &wi, float &pdf) {
    const v3 zero = mk3(0, 0, 0);
    if (b.kind == CL_DIFFUSE) { // material.h:76-85
        wi = cosine_hemisphere(ux, uy);
        if (!same_hemisphere(wo, wi)) wi.y = -wi.y;
        pdf = fabsf(wi.y) * AKB_INVPI;
        return scale3(AKB_INVPI, b.R);
    }
    if (b.kind == CL_GGX) { // material.h:122-137 + microfacet.h:123-148
        float phi = 2 * AKB_PI * uy;
        float t2 = b.alpha * b.alpha * ux / (1 - ux);
        float cosTheta = 1.0f / sqrtf(1 + t2);
        float sinTheta = sqrtf(max_std(0.0f, 1 - cosTheta * cosTheta));
        float sp, cp;
        sincos_ref(phi, sp, cp);
        v3 wh = mk3(cp * sinTheta, cosTheta, sp * sinTheta);
        if (!same_hemisphere(wo, wh)) wh = neg3(wh);
        wi = reflect3(wo, wh);
        if (!same_hemisphere(wo, wi)) {
            pdf = 0;
            return zero;
        }
        if (wh.y < 0) wh = neg3(wh);
        pdf = (ggx_D(b.alpha, wh) * fabsf(wh.y)) / (4.0f * fabsf(dot3(wo, wh)));
        return closure_evaluate(b, wo, wi);
    }
    if (b.kind == CL_DIELECTRIC) {
        // EXTENSION (not in the reference): smooth dielectric built from the reference's own
        // fr_dielectric / refract helpers. Reflect with probability F, else refract; no eta^2 radiance scaling.
        float F = fr_dielectric(wo.y, 1.0f, b.ior);
        if (ux < F) {
            wi = mk3(-wo.x, wo.y, -wo.z);
            pdf = F;
            float s = F / fabsf(wi.y);
            return mk3(b.R.x * s, b.R.y * s, b.R.z * s);
        }
        bool entering = wo.y > 0;
        float eta = entering ? 1.0f / b.ior : b.ior;
        v3 n = mk3(0, entering ? 1.0f : -1.0f, 0);
        if (!refract3(wo, n, eta, wi)) {
            pdf = 0;
            return zero;
        }
        pdf = 1.0f - F;
        float s = (1.0f - F) / fabsf(wi.y);
        return mk3(b.R.x * s, b.R.y * s, b.R.z * s);
    }
    pdf = 0;
    return zero;
}

// Material::select_material + get_bsdf (kernel/material.h:255-297). u = ctx.u1[0], which MaterialEvalContext draws
// from a COPY of the sampler (material.h:198-202), i.e. the same number the BSDF sample then uses as u.x.
// Returns the leaf material index; choice_pdf accumulates 1/frac or 1/(1-frac) (sic: a multiplication, :260-268).
// ImageTexture::evaluate (kernel/texture.h:45-49) + TImage::operator()(float2) (core/image.hpp:46-68):
// tc = fmod(texcoords, 1); tc.y = 1 - tc.y; texel int2(tc * resolution) (truncation), clamped to the image.
AKB_HD v3 tex_lookup(const SceneDev &S, int tex, float tu, float tv) {
    const int4 info = S.tex_info[tex];
    const int first = info.x, w = info.y, h = info.z;
    float x = fmodf(tu, 1.0f), y = fmodf(tv, 1.0f);
    y = 1.0f - y;
    int ix = (int)(x * (float)w), iy = (int)(y * (float)h);
    ix = ix < 0 ? 0 : (ix > w - 1 ? w - 1 : ix);
    iy = iy < 0 ? 0 : (iy > h - 1 ? h - 1 : iy);
    const float4 t = AKB_LD4(S.texels + (size_t)first + (size_t)ix + (size_t)iy * (size_t)w);
    return mk3(t.x, t.y, t.z);
}
// Triangle::texcoord(uv) = lerp3 of the corner texcoords (kernel/shape.h:40-42)
AKB_HD void tri_texcoord(const float4 &a, const float4 &b, float u, float v, float &tu, float &tv) {
    const v3 t = lerp3(mk3(a.x, a.y, 0.f), mk3(a.z, a.w, 0.f), mk3(b.x, b.y, 0.f), u, v);
    tu = t.x, tv = t.y;
}

AKB_HD int select_material(const SceneDev &S, int mat, float u, float tu, float tv, float &choice_pdf, bool tex_on = true) {
    const akb_material *mats = S.mats;
    choice_pdf = 1.0f;
    int guard = 0;
    while (mats[mat].type == AKB_MAT_MIX && guard++ < 64) {
        float frac = mats[mat].fraction;
        if (tex_on && S.n_textures && mats[mat].fraction_tex >= 0) frac = tex_lookup(S, mats[mat].fraction_tex, tu, tv).x;
        if (u < frac) {
            u = u / frac;
            choice_pdf *= 1.0f / frac;
            mat = mats[mat].child_b;
        } else {
            u = (u - frac) / (1.0f - frac);
            choice_pdf *= 1.0f / (1.0f - frac);
            mat = mats[mat].child_a;
        }
    }
    return mat;
}
AKB_HD int closure_kind_of(int type) {
    return type == AKB_MAT_DIFFUSE ? CL_DIFFUSE
                                   : (type == AKB_MAT_GLOSSY ? CL_GGX : (type == AKB_MAT_DIELECTRIC ? CL_DIELECTRIC : CL_NULL));
}
AKB_HD Bsdf material_get_bsdf(const SceneDev &S, int mat, float u, v3 ns, float tu, float tv, bool tex_on = true) {
    Bsdf b;
    int leaf = select_material(S, mat, u, tu, tv, b.choice_pdf, tex_on);
    const akb_material &m = S.mats[leaf];
    b.frame = frame_make(ns);
    b.R = mk3(m.color[0], m.color[1], m.color[2]);
    b.kind = closure_kind_of(m.type);
    if (tex_on && S.n_textures && m.color_tex >= 0 && b.kind != CL_NULL) b.R = tex_lookup(S, m.color_tex, tu, tv);
    b.alpha = 0.f, b.ior = m.ior;
    if (b.kind == CL_GGX) {
        float r = m.roughness;
        if (tex_on && S.n_textures && m.roughness_tex >= 0) r = tex_lookup(S, m.roughness_tex, tu, tv).x;
        r *= r; // material.h:225-226
        b.alpha = r;
    }
    return b;
}

// ---- lights: common/distribution.h:32-81, kernel/scene.h:79-90 ----------------------------------------------
AKB_HD int select_light(const SceneDev &S, float u, float &pdf) {
    if (S.n_lights == 0) {
        pdf = 0.0f;
        return -1;
    }
    int lo = 0, hi = S.n_lights + 1;
    const int last = hi;
    while (lo < hi) { // upper_bound: first index with cdf[idx] > u
        int mid = (lo + hi) / 2;
        if (S.light_cdf[mid] <= u)
            lo = mid + 1;
        else
            hi = mid;
    }
    int i = hi - 1;
    if (i < 0) i = 0;
    if (i > last - 2) i = last - 2;
    pdf = S.light_func[i] / (S.light_func_int * (float)S.n_lights);
    return i;
}

AKB_HD bool is_black(v3 c) { return !((c.x > 0.0f) || (c.y > 0.0f) || (c.z > 0.0f)); }

struct RaySeg {
    v3 o, d;
    float tmin, tmax;
};

// Live state of one camera path (kernel/pathtracer.h:53-58).
template <int SAMP> struct PathState {
    v3 L, beta;
    int depth;
    Rng<SAMP> rng;
};

enum { VTX_CONTINUE = 1, VTX_SHADOW = 2 };

// One iteration of run_megakernel's loop body AFTER scene.intersect() returned a hit (kernel/pathtracer.h:142-162):
// on_surface_scatter (:96-132), compute_direct_lighting (:65-91), throughput update; plus the two extensions
// (dielectric closure, Russian roulette). Returns VTX_* flags: CONTINUE -> `next` is the extension ray;
// SHADOW -> `shadow` must be tested with occlude() and `shadow_color` added to L when it is unoccluded.
template <int SAMP>
AKB_HD int path_vertex(const SceneDev &S, PathState<SAMP> &ps, v3 ray_d, float hu, float hv, int tri, int max_depth,
                       int rr_start, RaySeg &next, RaySeg &shadow, v3 &shadow_color, const bool tex_on = true) {
    // tex_on: a compile-time false in the kernel instantiation used for scenes without image textures (every BASELINE
    // config), so that none of the texture code is even present there
    const float4 *sr = S.shading + 6 * (size_t)tri;
    const float4 s0 = AKB_LD4(sr), s1 = AKB_LD4(sr + 1), s2 = AKB_LD4(sr + 2);
    const int mat = (int)akb_f2u(s0.w);
    if (mat < 0) return 0; // the reference dereferences a null material here; defined as "terminate"
    const akb_material &m = S.mats[mat];
    const v3 ng = mk3(s1.w, s2.w, AKB_LD4(sr + 3).w);
    const v3 wo = neg3(ray_d);
    float tu = 0.f, tv = 0.f; // si.texcoords (kernel/interaction.h:40-42); only needed when the scene has image textures
    if (tex_on && S.n_textures) tri_texcoord(AKB_LD4(S.tri_tc + 2 * (size_t)tri), AKB_LD4(S.tri_tc + 2 * (size_t)tri + 1), hu, hv, tu, tv);
    if (m.type == AKB_MAT_EMISSIVE) { // :102-111
        if (ps.depth == 0) {
            bool face_front = dot3(neg3(wo), ng) < 0.0f;
            if (m.double_sided || face_front) {
                v3 Le = mk3(m.color[0], m.color[1], m.color[2]);
                if (tex_on && S.n_textures && m.color_tex >= 0) Le = tex_lookup(S, m.color_tex, tu, tv);
                ps.L = add3(ps.L, mul3(ps.beta, Le));
            }
        }
        return 0;
    }
    if (!(ps.depth < max_depth)) return 0; // :112
    const float4 s3 = AKB_LD4(sr + 3), s4 = AKB_LD4(sr + 4), s5 = AKB_LD4(sr + 5);
    // SurfaceInteraction(uv, triangle), kernel/interaction.h:40-42
    const v3 p = lerp3(mk3(s0.x, s0.y, s0.z), mk3(s1.x, s1.y, s1.z), mk3(s2.x, s2.y, s2.z), hu, hv);
    const v3 ns = lerp3(mk3(s3.x, s3.y, s3.z), mk3(s4.x, s4.y, s4.z), mk3(s5.x, s5.y, s5.z), hu, hv);
    float bux, buy;
    ps.rng.next2d(bux, buy);
    const Bsdf bsdf = material_get_bsdf(S, mat, bux, ns, tu, tv, tex_on);
    const v3 wo_l = w2l(bsdf.frame, wo);
    v3 wi_l;
    float pdf;
    const v3 f = closure_sample(bsdf, bux, buy, wo_l, wi_l, pdf);
    const v3 wi = l2w(bsdf.frame, wi_l);
    pdf *= bsdf.choice_pdf;
    if (pdf == 0.0f) return 0; // :120-121
    const float angi = fabsf(dot3(ng, wi));
    next.o = p, next.d = wi, next.tmin = AKB_EPS / angi, next.tmax = INFINITY; // :126
    const v3 ev_beta = div3s(scale3(angi, f), pdf);                           // :127
    int flags = VTX_CONTINUE;
    // compute_direct_lighting(si, hit, select_light(scene)), :65-91,147-158
    float slx, sly;
    ps.rng.next2d(slx, sly);
    float light_pdf;
    const int li = select_light(S, slx, light_pdf);
    if (li >= 0) {
        float lux, luy;
        ps.rng.next2d(lux, luy);
        const float4 *lr = S.lights + 4 * (size_t)li;
        const float4 l0 = AKB_LD4(lr), l1 = AKB_LD4(lr + 1), l2 = AKB_LD4(lr + 2), l3 = AKB_LD4(lr + 3);
        // AreaLight::sample, kernel/light.h:58-71
        const float su0 = sqrtf(lux);
        const float b0 = 1 - su0, b1 = luy * su0;
        const v3 lp = lerp3(mk3(l0.x, l0.y, l0.z), mk3(l1.x, l1.y, l1.z), mk3(l2.x, l2.y, l2.z), b0, b1);
        const v3 lng = mk3(l3.x, l3.y, l3.z);
        v3 lwi = sub3(lp, p);
        const float dist_sqr = dot3(lwi, lwi);
        lwi = div3s(lwi, sqrtf(dist_sqr));
        const float area = l0.w;
        const float lpdf = dist_sqr / max_std(0.0f, -dot3(lwi, lng)) / area;
        if (!(lpdf <= 0.0f)) { // :76
            light_pdf *= lpdf;
            const v3 fe = closure_evaluate(bsdf, w2l(bsdf.frame, wo), w2l(bsdf.frame, lwi));
            v3 Le = mk3(l1.w, l2.w, l3.w);
            if (tex_on && S.n_textures) { // sample.L = color->evaluate(triangle.texcoord(coords)), kernel/light.h:65
                const float4 t0 = AKB_LD4(S.light_tc + 2 * (size_t)li), t1 = AKB_LD4(S.light_tc + 2 * (size_t)li + 1);
                const int ltex = (int)akb_f2u(t1.z);
                if (ltex >= 0) {
                    float lu, lv;
                    tri_texcoord(t0, t1, b0, b1, lu, lv);
                    Le = tex_lookup(S, ltex, lu, lv);
                }
            }
            const v3 ff = scale3(fabsf(dot3(ns, lwi)), mul3(Le, fe));
            const v3 color = div3s(mul3(ps.beta, ff), light_pdf);
            if (!is_black(color)) {
                shadow.o = lp, shadow.d = neg3(lwi);
                shadow.tmin = AKB_EPS / fabsf(dot3(lwi, lng));
                shadow.tmax = sqrtf(dist_sqr) * (1.0f - AKB_SHADOW_EPS);
                shadow_color = color;
                flags |= VTX_SHADOW;
            }
        }
    }
    ps.beta = mul3(ps.beta, ev_beta); // :159-162
    ps.depth++;
    if (rr_start > 0 && ps.depth >= rr_start) {
        // EXTENSION: Russian roulette. q = min(max(beta), 0.95); one next1d() draw.
        float q = min_std(max_std(max_std(ps.beta.x, ps.beta.y), ps.beta.z), 0.95f);
        float u = ps.rng.next1d();
        if (!(u < q)) return flags & ~VTX_CONTINUE;
        ps.beta = div3s(ps.beta, q);
    }
    return flags;
}

// Tile::add_sample (core/film.h:66-70) + the GPU integrator's optional per-sample clamp
// (gpu/cuda/integrator.cpp:196-198: clamp_zero then min(L, ray_clamp)).
AKB_HD v3 clamp_sample(v3 L, float ray_clamp) {
    if (ray_clamp > 0.0f) {
        float *c = &L.x;
        for (int k = 0; k < 3; k++) {
            float v = c[k];
            v = (v != v) ? 0.0f : max_std(0.0f, v);
            c[k] = min_std(v, ray_clamp);
        }
    }
    return L;
}

// The secondary ray of AmbientOcclusion's Li (kernel/integrators/cpu/integrator.cpp:46-51,
// gpu/cuda/integrator.cpp:49-54): Frame3f(trig.ng()), one cosine-weighted direction, origin trig.p(uv), Ray ctor
// defaults (tmin = Eps, tmax = inf). "closest hit nearer than `occlude`" is asked as an any-hit query with tmax = occlude.
template <int SAMP> AKB_HD RaySeg ao_ray(const SceneDev &S, Rng<SAMP> &rng, float hu, float hv, int tri, float occlude) {
    const float4 *sr = S.shading + 6 * (size_t)tri;
    const float4 s0 = AKB_LD4(sr), s1 = AKB_LD4(sr + 1), s2 = AKB_LD4(sr + 2);
    const v3 ng = mk3(s1.w, s2.w, AKB_LD4(sr + 3).w);
    const Frame fr = frame_make(ng);
    float ux, uy;
    rng.next2d(ux, uy);
    RaySeg r;
    r.d = l2w(fr, cosine_hemisphere(ux, uy));
    r.o = lerp3(mk3(s0.x, s0.y, s0.z), mk3(s1.x, s1.y, s1.z), mk3(s2.x, s2.y, s2.z), hu, hv);
    r.tmin = AKB_EPS, r.tmax = occlude;
    return r;
}

} // namespace akb
