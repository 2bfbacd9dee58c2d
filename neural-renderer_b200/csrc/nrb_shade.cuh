// nrb_shade.cuh — environment lookup, tangent frame, ray generation and the Disney BRDF.
//
// Arithmetic follows the reference expression by expression (operation order matters because the
// parity tests compare against the reference headers compiled for the host):
//   generatePrimaryRay            src/mesh_traversal.cuh:40-55
//   sampleEnvironment(+WithClamp) src/cuda_renderer_neural.cu:115-171, 193-202
//   buildTangentSpace             src/cuda_renderer_neural.cu:174-190
//   Disney lobes / sampling / pdf src/disney_brdf.cuh:31-217, 224-286, 395-508
// including its quirks: sin_theta = 1 - cos^2 (no sqrt) in the GGX and GTR1 samplers, the GGX
// sampler's own frame built from world z, and the un-clamped specular pdf.
#pragma once

#include "nrb_scene.cuh"

namespace nrb {

constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kTwoPi = 2.0f * 3.14159265358979323846f;

struct Bsdf {
    V3 base;
    float metallic, roughness, specular, specular_tint, anisotropy;
    float sheen, sheen_tint, clearcoat, clearcoat_gloss, ior, spec_trans;
};

__device__ __forceinline__ Bsdf bsdf_from(const MaterialDev& g, V3 albedo, V3 mp) {
    Bsdf b;
    b.base = albedo;
    b.metallic = mp.x; b.roughness = mp.y; b.specular = mp.z;
    b.specular_tint = g.specular_tint.value; b.anisotropy = g.anisotropy.value;
    b.sheen = g.sheen.value; b.sheen_tint = g.sheen_tint.value;
    b.clearcoat = g.clearcoat.value; b.clearcoat_gloss = g.clearcoat_gloss.value;
    b.ior = g.ior; b.spec_trans = g.specular_transmission;
    return b;
}

// Pixel of a local (tile-owned) row -> global row.  world == 1 is the identity.
__device__ __forceinline__ int global_row(const FrameParams& fp, int localRow) {
    if (fp.world <= 1) return localRow;
    const int band = localRow / fp.bandRows;
    return (band * fp.world + fp.rank) * fp.bandRows + (localRow - band * fp.bandRows);
}

__device__ __forceinline__ V3 primary_direction(const FrameParams& fp, int x, int y, uint32_t& rng) {
    const float jx = rng_next(rng), jy = rng_next(rng);
    float u = (static_cast<float>(x) + jx) / static_cast<float>(fp.width);
    float v = 1.0f - (static_cast<float>(y) + jy) / static_cast<float>(fp.height);
    u = u * 2.0f - 1.0f;
    v = v * 2.0f - 1.0f;
    const V3 d = fp.camForward + fp.camRight * (u * fp.aspect * fp.tanHalfFov) + fp.camUp * (v * fp.tanHalfFov);
    return normalize(d);
}

__device__ inline V3 env_lookup(const EnvDev& env, V3 dir) {
    if (env.pixels && env.width > 0 && env.height > 0) {
        if (env.rotation != 0.0f) {
            const float ang = env.rotation * (3.14159265358979323846f / 180.0f);
            const float ca = cosf(ang), sa = sinf(ang);
            const float nx = dir.x * ca + dir.z * sa;
            const float nz = -dir.x * sa + dir.z * ca;
            dir.x = nx;
            dir.z = nz;
        }
        const float kIp = 0.3183098861837907f;
        float u = atan2f(dir.x, -dir.z) * kIp * 0.5f + 0.5f;
        float v = acosf(clampf(dir.y, -1.0f, 1.0f)) * kIp;
        u = clampf(u, 0.0f, 1.0f);
        v = clampf(v, 0.0f, 1.0f);
        const float x = u * static_cast<float>(env.width) - 0.5f;
        const float y = v * static_cast<float>(env.height) - 0.5f;
        int x0 = static_cast<int>(floorf(x)), y0 = static_cast<int>(floorf(y));
        int x1 = x0 + 1, y1 = y0 + 1;
        x0 = max(0, min(x0, env.width - 1));
        y0 = max(0, min(y0, env.height - 1));
        x1 = max(0, min(x1, env.width - 1));
        y1 = max(0, min(y1, env.height - 1));
        const float tx = x - static_cast<float>(x0), ty = y - static_cast<float>(y0);
        const V3 c00 = ld3(env.pixels, y0 * env.width + x0), c10 = ld3(env.pixels, y0 * env.width + x1);
        const V3 c01 = ld3(env.pixels, y1 * env.width + x0), c11 = ld3(env.pixels, y1 * env.width + x1);
        const V3 c0 = c00 * (1.0f - tx) + c10 * tx;
        const V3 c1 = c01 * (1.0f - tx) + c11 * tx;
        const V3 e = (c0 * (1.0f - ty) + c1 * ty) * env.strength;
        return v3(fminf(e.x, 100.0f), fminf(e.y, 100.0f), fminf(e.z, 100.0f));
    }
    const float t = 0.5f * (dir.y + 1.0f);
    return mix3(v3s(1.0f), v3(0.5f, 0.7f, 1.0f), t);
}

__device__ __forceinline__ V3 env_clamped(const EnvDev& env, V3 dir, float cap) {
    const V3 e = env_lookup(env, dir);
    return v3(fminf(e.x, cap), fminf(e.y, cap), fminf(e.z, cap));
}

__device__ __forceinline__ void tangent_frame(V3 n, V3& t, V3& b) {
    V3 axis = v3s(0.0f);
    if (n.x < 0.6f && n.x > -0.6f) axis.x = 1.0f;
    else if (n.y < 0.6f && n.y > -0.6f) axis.y = 1.0f;
    else if (n.z < 0.6f && n.z > -0.6f) axis.z = 1.0f;
    else axis.x = 1.0f;
    t = normalize(cross(axis, n));
    b = normalize(cross(n, t));
}

// ---- Disney terms --------------------------------------------------------------------------

__device__ __forceinline__ float sq(float x) { return x * x; }
__device__ __forceinline__ float schlick5(float c) {
    const float m = clamp01(1.0f - c);
    const float m2 = m * m;
    return m2 * m2 * m;
}
__device__ __forceinline__ float lum(V3 c) { return 0.2126f * c.x + 0.7152f * c.y + 0.0722f * c.z; }
__device__ __forceinline__ float d_gtr1(float ch, float a) {
    if (a >= 1.0f) return kInvPi;
    const float a2 = a * a;
    return kInvPi * (a2 - 1.0f) / (logf(a2) * (1.0f + (a2 - 1.0f) * ch * ch));
}
__device__ __forceinline__ float d_gtr2(float ch, float a) {
    const float a2 = a * a;
    return kInvPi * a2 / sq(1.0f + (a2 - 1.0f) * ch * ch);
}
__device__ __forceinline__ float d_gtr2_aniso(float hn, float hx, float hy, V2 a) {
    return kInvPi / (a.x * a.y * sq(sq(hx / a.x) + sq(hy / a.y) + hn * hn));
}
__device__ __forceinline__ float g_smith(float no, float ag) {
    const float a = ag * ag, b = no * no;
    return 1.0f / (no + sqrtf(a + b - a * b));
}
__device__ __forceinline__ float g_smith_aniso(float no, float ox, float oy, V2 a) {
    return 1.0f / (no + sqrtf(sq(ox * a.x) + sq(oy * a.y) + sq(no)));
}
__device__ __forceinline__ bool same_side(V3 wo, V3 wi, V3 n) { return dot(wo, n) * dot(wi, n) > 0.0f; }
__device__ __forceinline__ V3 tint_of(const Bsdf& m) {
    const float l = lum(m.base);
    return l > 0.0f ? m.base / l : v3s(1.0f);
}
__device__ __forceinline__ float fresnel_diel(float ci, float ei, float et) {
    float g = sq(et) / sq(ei) - 1.0f + sq(ci);
    if (g < 0.0f) return 1.0f;
    g = sqrtf(g);
    const float a = (g - ci) / (g + ci);
    const float b = (ci * (g + ci) - 1.0f) / (ci * (g - ci) + 1.0f);
    return 0.5f * a * a * (1.0f + b * b);
}

__device__ inline V3 disney_f(const Bsdf& m, V3 n, V3 wo, V3 wi, V3 tg, V3 bt) {
    if (!same_side(wo, wi, n)) {
        if (m.spec_trans > 0.0f) {
            const float on = dot(wo, n), in = dot(wi, n);
            V3 st = v3s(0.0f);
            if (!(on == 0.0f || in == 0.0f)) {
                const bool entering = on > 0.0f;
                const float eo = entering ? 1.0f : m.ior, ei = entering ? m.ior : 1.0f;
                const V3 h = normalize(wo + wi * ei / eo);
                const float a = fmaxf(0.001f, m.roughness * m.roughness);
                const float d = d_gtr2(fabsf(dot(n, h)), a);
                const float f = fresnel_diel(fabsf(dot(wi, n)), eo, ei);
                const float g = g_smith(fabsf(dot(n, wi)), a) * g_smith(fabsf(dot(n, wo)), a);
                const float ih = dot(wi, h), oh = dot(wo, h);
                const float c = fabsf(oh) / fabsf(dot(wo, n)) * fabsf(ih) / fabsf(dot(wi, n)) * sq(eo) /
                                sq(eo * oh + ei * ih);
                st = m.base * (c * (1.0f - f) * g * d);
            }
            return st * ((1.0f - m.metallic) * m.spec_trans);
        }
        return v3s(0.0f);
    }
    const V3 h = normalize(wi + wo);
    // clearcoat
    const float ca = mixf(0.1f, 0.001f, m.clearcoat_gloss);
    const float coat = 0.25f * m.clearcoat * d_gtr1(dot(n, h), ca) * mixf(0.04f, 1.0f, schlick5(dot(wi, n))) *
                       (g_smith(dot(n, wi), 0.25f) * g_smith(dot(n, wo), 0.25f));
    // sheen
    const V3 tint = tint_of(m);
    const V3 sheen = mix3(v3s(1.0f), tint, m.sheen_tint) * (schlick5(dot(wi, h)) * m.sheen);
    // diffuse
    const float no = fabsf(dot(wo, n)), ni = fabsf(dot(wi, n));
    const float ih = dot(wi, h);
    const float fd90 = 0.5f + 2.0f * m.roughness * ih * ih;
    const V3 diffuse = m.base * kInvPi * mixf(1.0f, fd90, schlick5(ni)) * mixf(1.0f, fd90, schlick5(no));
    // gloss
    const V3 spec = mix3(m.specular * 0.08f * mix3(v3s(1.0f), tint, m.specular_tint), m.base, m.metallic);
    const V3 fr = mix3(spec, v3s(1.0f), schlick5(dot(wi, h)));
    V3 gloss;
    if (m.anisotropy == 0.0f) {
        const float a = fmaxf(0.001f, m.roughness * m.roughness);
        const float d = d_gtr2(dot(n, h), a);
        const float g = g_smith(dot(n, wi), a) * g_smith(dot(n, wo), a);
        gloss = fr * (d * g);
    } else {
        const float asp = sqrtf(1.0f - m.anisotropy * 0.9f);
        const float a = m.roughness * m.roughness;
        const V2 al = V2{fmaxf(0.001f, a / asp), fmaxf(0.001f, a * asp)};
        const float d = d_gtr2_aniso(dot(n, h), fabsf(dot(h, tg)), fabsf(dot(h, bt)), al);
        const float g = g_smith_aniso(dot(n, wi), fabsf(dot(wi, tg)), fabsf(dot(wi, bt)), al) *
                        g_smith_aniso(dot(n, wo), fabsf(dot(wo, tg)), fabsf(dot(wo, bt)), al);
        gloss = fr * (d * g);
    }
    return (diffuse + sheen) * ((1.0f - m.metallic) * (1.0f - m.spec_trans)) + gloss + v3s(coat);
}

__device__ inline float disney_pdf3(const Bsdf& m, V3 n, V3 wo, V3 wi, V3 tg, V3 bt) {
    if (!same_side(wo, wi, n)) return 0.0f;
    const float a = fmaxf(0.001f, m.roughness * m.roughness);
    const float asp = sqrtf(1.0f - m.anisotropy * 0.9f);
    const V2 al = V2{fmaxf(0.001f, a / asp), fmaxf(0.001f, a * asp)};
    const float ca = mixf(0.1f, 0.001f, m.clearcoat_gloss);
    const float pdfD = fmaxf(0.0f, dot(n, wi)) * kInvPi;
    const V3 h = normalize(wi + wo);
    float pdfS;
    if (m.anisotropy == 0.0f) {
        const float nh = dot(n, h), ho = dot(h, wo);
        pdfS = d_gtr2(nh, a) * nh / (4.0f * ho);
    } else {
        const float ch = fabsf(dot(n, h));
        pdfS = d_gtr2_aniso(ch, fabsf(dot(h, tg)), fabsf(dot(h, bt)), al) * ch / (4.0f * fabsf(dot(wo, h)));
    }
    const float chc = dot(n, h);
    const float pdfC = d_gtr1(chc, ca) * chc / (4.0f * dot(wo, h));
    return (pdfD + pdfS + pdfC) / 3.0f;
}

// Local-frame half/direction builders of the reference samplers.
__device__ __forceinline__ V3 around_z_frame(V3 n, float hx, float hy, float hz) {
    const V3 up = fabsf(n.z) < 0.999f ? v3(0.0f, 0.0f, 1.0f) : v3(1.0f, 0.0f, 0.0f);
    const V3 t = normalize(cross(up, n));
    const V3 b = cross(n, t);
    return normalize(t * hx + b * hy + n * hz);
}

__device__ inline V3 disney_sample3(const Bsdf& m, V3 n, V3 wo, V3 tg, V3 bt, float u1, float u2, float u3,
                                    float& pdf) {
    if (m.spec_trans >= 1.0f) { pdf = 0.0f; return v3s(0.0f); }
    int lobe = static_cast<int>(u3 * 3.0f);
    lobe = lobe > 2 ? 2 : lobe;
    V3 wi;
    if (lobe == 0) {
        const float phi = kTwoPi * u1;
        const float ct = sqrtf(u2), st = sqrtf(1.0f - u2);
        wi = around_z_frame(n, st * cosf(phi), st * sinf(phi), ct);
    } else {
        V3 h;
        if (lobe == 1) {
            const float a = fmaxf(0.001f, m.roughness * m.roughness);
            if (m.anisotropy == 0.0f) {
                const float phi = kTwoPi * u1;
                const float ct = sqrtf((1.0f - u2) / (1.0f + (a * a - 1.0f) * u2));
                const float st = 1.0f - ct * ct;
                h = around_z_frame(n, st * cosf(phi), st * sinf(phi), ct);
            } else {
                const float asp = sqrtf(1.0f - m.anisotropy * 0.9f);
                const V2 al = V2{fmaxf(0.001f, a / asp), fmaxf(0.001f, a * asp)};
                const float x = kTwoPi * u1;
                h = normalize(tg * (al.x * cosf(x) * sqrtf(u2 / (1.0f - u2))) +
                              bt * (al.y * sinf(x) * sqrtf(u2 / (1.0f - u2))) + n);
            }
        } else {
            const float a = mixf(0.1f, 0.001f, m.clearcoat_gloss);
            const float phi = kTwoPi * u1;
            const float a2 = a * a;
            const float c2 = (1.0f - powf(a2, 1.0f - u2)) / (1.0f - a2);
            const float ct = sqrtf(c2), st = 1.0f - c2;
            h = normalize(tg * (st * cosf(phi)) + bt * (st * sinf(phi)) + n * ct);
        }
        const V3 in = -wo;
        wi = in - h * (2.0f * dot(in, h));
        if (!same_side(wo, wi, n)) { pdf = 0.0f; return wi; }
    }
    pdf = disney_pdf3(m, n, wo, wi, tg, bt);
    return wi;
}

}  // namespace nrb
