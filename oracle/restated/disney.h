// oracle/restated/disney.h — TEST INFRASTRUCTURE (CPU oracle "port" build), never part of the product.
//
// Plain C++ restatement of the reference's Disney BRDF (src/disney_brdf.cuh): the terms used by
// disney_eval (:88-217), the half-vector samplers (:224-286) and the uniform three-lobe sampler with
// its pdf (:395-508).  Upstream quirks are kept on purpose and marked QUIRK.  Pinned against the
// reference header itself by tests/test_oracle_pinning.py wherever /root/reference exists.
#pragma once

#include "geometry.h"

namespace disney_detail {
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kPi2 = 2.0f * 3.14159265358979323846f;
inline float sq(float v) { return v * v; }
inline float mixf(float a, float b, float t) { return a * (1.0f - t) + b * t; }
inline float pow5(float c) {  // schlick_weight :31-35
    const float m = fminf(fmaxf(1.0f - c, 0.0f), 1.0f);
    const float m2 = m * m;
    return m2 * m2 * m;
}
inline float gtr1(float ch, float a) {  // :49-55
    if (a >= 1.0f) return kInvPi;
    const float a2 = a * a;
    return kInvPi * (a2 - 1.0f) / (logf(a2) * (1.0f + (a2 - 1.0f) * ch * ch));
}
inline float gtr2(float ch, float a) {  // :57-60
    const float a2 = a * a;
    return kInvPi * a2 / sq(1.0f + (a2 - 1.0f) * ch * ch);
}
inline float gtr2Aniso(float hn, float hx, float hy, Vec2 a) {  // :62-64
    return kInvPi / (a.x * a.y * sq(sq(hx / a.x) + sq(hy / a.y) + hn * hn));
}
inline float smithG(float no, float ag) {  // :66-70
    const float a = ag * ag, b = no * no;
    return 1.0f / (no + sqrtf(a + b - a * b));
}
inline float smithGAniso(float no, float ox, float oy, Vec2 a) {  // :72-74
    return 1.0f / (no + sqrtf(sq(ox * a.x) + sq(oy * a.y) + sq(no)));
}
inline bool sameSide(Vec3 wo, Vec3 wi, Vec3 n) { return dot(wo, n) * dot(wi, n) > 0.0f; }
inline Vec3 tintOf(Vec3 base) {
    const float l = 0.2126f * base.x + 0.7152f * base.y + 0.0722f * base.z;
    return l > 0.0f ? base / l : Vec3(1.0f, 1.0f, 1.0f);
}
inline float fresnelDielectric(float ci, float ei, float et) {  // :37-47
    float g = sq(et) / sq(ei) - 1.0f + sq(ci);
    if (g < 0.0f) return 1.0f;
    g = sqrtf(g);
    const float a = (g - ci) / (g + ci);
    const float b = (ci * (g + ci) - 1.0f) / (ci * (g - ci) + 1.0f);
    return 0.5f * a * a * (1.0f + b * b);
}
inline Vec2 anisoAlpha(float roughnessSq, float anisotropy) {
    const float aspect = sqrtf(1.0f - anisotropy * 0.9f);
    return Vec2(fmaxf(0.001f, roughnessSq / aspect), fmaxf(0.001f, roughnessSq * aspect));
}
// Direction (hx, hy, hz) expressed in the frame the upstream samplers build from world z (:232-236).
inline Vec3 inWorldZFrame(Vec3 n, float hx, float hy, float hz) {
    const Vec3 up = fabsf(n.z) < 0.999f ? Vec3(0.0f, 0.0f, 1.0f) : Vec3(1.0f, 0.0f, 0.0f);
    const Vec3 t = normalize(cross(up, n));
    const Vec3 b = cross(n, t);
    return normalize(t * hx + b * hy + n * hz);
}
}  // namespace disney_detail

inline float saturate(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }

// disney_eval :193-217 with the lobes :88-190 folded in.
inline Vec3 disney_eval(const Material& mat, Vec3 n, Vec3 wo, Vec3 wi, Vec3 tangent, Vec3 bitangent) {
    using namespace disney_detail;
    const Vec3 base = mat.base_color.value;
    const float rough = mat.roughness.value, metallic = mat.metallic.value;
    if (!sameSide(wo, wi, n)) {
        if (!(mat.specular_transmission > 0.0f)) return Vec3();
        const float on = dot(wo, n), in = dot(wi, n);
        Vec3 st;
        if (on != 0.0f && in != 0.0f) {  // disney_microfacet_transmission_isotropic :163-190
            const bool entering = on > 0.0f;
            const float eo = entering ? 1.0f : mat.ior, ei = entering ? mat.ior : 1.0f;
            const Vec3 h = normalize(wo + wi * ei / eo);
            const float a = fmaxf(0.001f, rough * rough);
            const float d = gtr2(fabsf(dot(n, h)), a);
            const float f = fresnelDielectric(fabsf(dot(wi, n)), eo, ei);
            const float g = smithG(fabsf(dot(n, wi)), a) * smithG(fabsf(dot(n, wo)), a);
            const float ih = dot(wi, h), oh = dot(wo, h);
            const float c = fabsf(oh) / fabsf(dot(wo, n)) * fabsf(ih) / fabsf(dot(wi, n)) * sq(eo) / sq(eo * oh + ei * ih);
            st = base * (c * (1.0f - f) * g * d);
        }
        return st * ((1.0f - metallic) * mat.specular_transmission);
    }
    const Vec3 h = normalize(wi + wo);
    const Vec3 tint = tintOf(base);
    // clear coat :111-120
    const float coatAlpha = mixf(0.1f, 0.001f, mat.clearcoat_gloss.value);
    const float coat = 0.25f * mat.clearcoat.value * gtr1(dot(n, h), coatAlpha) * mixf(0.04f, 1.0f, pow5(dot(wi, n))) *
                       (smithG(dot(n, wi), 0.25f) * smithG(dot(n, wo), 0.25f));
    // sheen :99-109
    const Vec3 sheen = lerp(Vec3(1.0f, 1.0f, 1.0f), tint, mat.sheen_tint.value) * (pow5(dot(wi, h)) * mat.sheen.value);
    // diffuse :84-97
    const float ih = dot(wi, h);
    const float fd90 = 0.5f + 2.0f * rough * ih * ih;
    const Vec3 diffuse = base * kInvPi * mixf(1.0f, fd90, pow5(fabsf(dot(wi, n)))) * mixf(1.0f, fd90, pow5(fabsf(dot(wo, n))));
    // gloss :122-161
    const Vec3 spec0 = lerp(mat.specular.value * 0.08f * lerp(Vec3(1.0f, 1.0f, 1.0f), tint, mat.specular_tint.value), base, metallic);
    const Vec3 fres = lerp(spec0, Vec3(1.0f, 1.0f, 1.0f), pow5(dot(wi, h)));
    Vec3 gloss;
    if (mat.anisotropy.value == 0.0f) {
        const float a = fmaxf(0.001f, rough * rough);
        gloss = fres * (gtr2(dot(n, h), a) * (smithG(dot(n, wi), a) * smithG(dot(n, wo), a)));
    } else {
        const Vec2 al = anisoAlpha(rough * rough, mat.anisotropy.value);
        const float d = gtr2Aniso(dot(n, h), fabsf(dot(h, tangent)), fabsf(dot(h, bitangent)), al);
        const float g = smithGAniso(dot(n, wi), fabsf(dot(wi, tangent)), fabsf(dot(wi, bitangent)), al) *
                        smithGAniso(dot(n, wo), fabsf(dot(wo, tangent)), fabsf(dot(wo, bitangent)), al);
        gloss = fres * (d * g);
    }
    return (diffuse + sheen) * ((1.0f - metallic) * (1.0f - mat.specular_transmission)) + gloss + Vec3(coat, coat, coat);
}

// disney_pdf_3component :468-508 — mean of the three lobe pdfs.  QUIRK: the specular pdf is not clamped.
inline float disney_pdf_3component(const Material& mat, Vec3 n, Vec3 wo, Vec3 wi, Vec3 tangent, Vec3 bitangent) {
    using namespace disney_detail;
    if (!sameSide(wo, wi, n)) return 0.0f;
    const float a = fmaxf(0.001f, mat.roughness.value * mat.roughness.value);
    const Vec2 al = anisoAlpha(a, mat.anisotropy.value);
    const float coatAlpha = mixf(0.1f, 0.001f, mat.clearcoat_gloss.value);
    const float pd = fmaxf(0.0f, dot(n, wi)) * kInvPi;
    const Vec3 h = normalize(wi + wo);
    float ps;
    if (mat.anisotropy.value == 0.0f) {
        const float nh = dot(n, h);
        ps = gtr2(nh, a) * nh / (4.0f * dot(h, wo));
    } else {  // gtr_2_aniso_pdf :306-317
        const float ch = fabsf(dot(n, h));
        ps = gtr2Aniso(ch, fabsf(dot(h, tangent)), fabsf(dot(h, bitangent)), al) * ch / (4.0f * fabsf(dot(wo, h)));
    }
    const float nh = dot(n, h);  // gtr_1_pdf :294-304
    const float pc = gtr1(nh, coatAlpha) * nh / (4.0f * dot(wo, h));
    return (pd + ps + pc) / 3.0f;
}

// disney_sample_3component :395-465 — lobe = floor(3 u3): cosine hemisphere / GGX (iso or aniso) / clearcoat GTR1.
// QUIRK: sin(theta) = 1 - cos^2(theta) without the square root in the GGX and GTR1 samplers (:227, :266).
inline Vec3 disney_sample_3component(const Material& mat, Vec3 n, Vec3 wo, Vec3 tangent, Vec3 bitangent, float u1, float u2,
                                     float u3, float* pdfOut) {
    using namespace disney_detail;
    auto give = [&](float p) { if (pdfOut) *pdfOut = p; };
    if (mat.specular_transmission >= 1.0f) {
        give(0.0f);
        return Vec3();
    }
    int lobe = static_cast<int>(u3 * 3.0f);
    if (lobe > 2) lobe = 2;
    const float phi = kPi2 * u1;
    Vec3 wi;
    if (lobe == 0) {  // sample_cosine_hemisphere :240-253
        const float ct = sqrtf(u2), st = sqrtf(1.0f - u2);
        wi = inWorldZFrame(n, st * cosf(phi), st * sinf(phi), ct);
    } else {
        Vec3 h;
        if (lobe == 1) {
            const float a = fmaxf(0.001f, mat.roughness.value * mat.roughness.value);
            if (mat.anisotropy.value == 0.0f) {  // sample_ggx :224-238
                const float ct = sqrtf((1.0f - u2) / (1.0f + (a * a - 1.0f) * u2));
                const float st = 1.0f - ct * ct;
                h = inWorldZFrame(n, st * cosf(phi), st * sinf(phi), ct);
            } else {  // sample_gtr_2_aniso_h :275-286
                const Vec2 al = anisoAlpha(a, mat.anisotropy.value);
                h = normalize(tangent * (al.x * cosf(phi) * sqrtf(u2 / (1.0f - u2))) +
                              bitangent * (al.y * sinf(phi) * sqrtf(u2 / (1.0f - u2))) + n);
            }
        } else {  // sample_gtr_1_h :255-273
            const float a = mixf(0.1f, 0.001f, mat.clearcoat_gloss.value);
            const float a2 = a * a;
            const float c2 = (1.0f - powf(a2, 1.0f - u2)) / (1.0f - a2);
            const float st = 1.0f - c2;
            h = normalize(tangent * (st * cosf(phi)) + bitangent * (st * sinf(phi)) + n * sqrtf(c2));
        }
        const Vec3 in = -wo;  // reflect(-wo, h) :289-291
        wi = in - h * (2.0f * dot(in, h));
        if (!sameSide(wo, wi, n)) {
            give(0.0f);
            return wi;
        }
    }
    give(disney_pdf_3component(mat, n, wo, wi, tangent, bitangent));
    return wi;
}
