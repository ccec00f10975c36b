// Device arithmetic for the physics tick.  FP32 throughout, compiled with
// -fmad=false -prec-div=true -prec-sqrt=true so that every operation rounds like the
// reference's scalar x86-64 build (no FMA contraction).  Operation ORDER follows glm
// 0.9.9.9 as vendored by the reference (external/glm); each helper cites the glm or
// reference line it mirrors.  No tensor cores: nothing here is a dense contraction.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace icx {

struct V3 { float x, y, z; };
struct Q4 { float w, x, y, z; };
struct M3 { V3 c[3]; };          // column-major
struct M4 { V3 c[3]; V3 t; };    // affine [c0 c1 c2 | t]

#define ICX_DI __host__ __device__ __forceinline__
#define ICX_DN static __host__ __device__ __noinline__

ICX_DI V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
ICX_DI V3 v3(float4 f) { return v3(f.x, f.y, f.z); }
ICX_DI V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
ICX_DI V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
ICX_DI V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
ICX_DI V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
ICX_DI V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
ICX_DI V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
ICX_DI V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
// glm/detail/func_geometric.inl:48-55, :68-79, :82-90; func_exponential.inl:134-139
ICX_DI float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
ICX_DI V3 cross(V3 x, V3 y) { return v3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
ICX_DI float length2(V3 a) { return dot(a, a); }
ICX_DI float length(V3 a) { return sqrtf(dot(a, a)); }
ICX_DI V3 normalize(V3 a) { return a * (1.0f / sqrtf(dot(a, a))); }
// std::min / std::max argument conventions (they differ from fminf/fmaxf on NaN)
ICX_DI float smin(float a, float b) { return (b < a) ? b : a; }
ICX_DI float smax(float a, float b) { return (a < b) ? b : a; }

// glm mat3*vec3, detail/type_mat3x3.inl:468-474
ICX_DI V3 mul(const M3& m, V3 v) {
    return v3(m.c[0].x * v.x + m.c[1].x * v.y + m.c[2].x * v.z,
              m.c[0].y * v.x + m.c[1].y * v.y + m.c[2].y * v.z,
              m.c[0].z * v.x + m.c[1].z * v.y + m.c[2].z * v.z);
}
// glm mat3*mat3, detail/type_mat3x3.inl:486-519
ICX_DI M3 mul(const M3& a, const M3& b) {
    M3 r;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        V3 bc = b.c[c];
        r.c[c] = v3(a.c[0].x * bc.x + a.c[1].x * bc.y + a.c[2].x * bc.z,
                    a.c[0].y * bc.x + a.c[1].y * bc.y + a.c[2].y * bc.z,
                    a.c[0].z * bc.x + a.c[1].z * bc.y + a.c[2].z * bc.z);
    }
    return r;
}
ICX_DI M3 mul(const M3& a, float s) { M3 r; r.c[0] = a.c[0] * s; r.c[1] = a.c[1] * s; r.c[2] = a.c[2] * s; return r; }
ICX_DI M3 transpose(const M3& m) {
    M3 r;
    r.c[0] = v3(m.c[0].x, m.c[1].x, m.c[2].x);
    r.c[1] = v3(m.c[0].y, m.c[1].y, m.c[2].y);
    r.c[2] = v3(m.c[0].z, m.c[1].z, m.c[2].z);
    return r;
}
// glm::inverse(mat3), detail/func_matrix.inl:322-344
ICX_DI M3 inverse(const M3& m) {
    const float m00 = m.c[0].x, m01 = m.c[0].y, m02 = m.c[0].z;
    const float m10 = m.c[1].x, m11 = m.c[1].y, m12 = m.c[1].z;
    const float m20 = m.c[2].x, m21 = m.c[2].y, m22 = m.c[2].z;
    float ood = 1.0f / (+m00 * (m11 * m22 - m21 * m12) - m10 * (m01 * m22 - m21 * m02) + m20 * (m01 * m12 - m11 * m02));
    M3 r;
    r.c[0].x = +(m11 * m22 - m21 * m12) * ood;
    r.c[1].x = -(m10 * m22 - m20 * m12) * ood;
    r.c[2].x = +(m10 * m21 - m20 * m11) * ood;
    r.c[0].y = -(m01 * m22 - m21 * m02) * ood;
    r.c[1].y = +(m00 * m22 - m20 * m02) * ood;
    r.c[2].y = -(m00 * m21 - m20 * m01) * ood;
    r.c[0].z = +(m01 * m12 - m11 * m02) * ood;
    r.c[1].z = -(m00 * m12 - m10 * m02) * ood;
    r.c[2].z = +(m00 * m11 - m10 * m01) * ood;
    return r;
}
ICX_DI M3 mat3_of(const M4& m) { M3 r; r.c[0] = m.c[0]; r.c[1] = m.c[1]; r.c[2] = m.c[2]; return r; }
ICX_DI M3 diag3(V3 d) { M3 m; m.c[0] = v3(d.x, 0.0f, 0.0f); m.c[1] = v3(0.0f, d.y, 0.0f); m.c[2] = v3(0.0f, 0.0f, d.z); return m; }
// glm mat4*vec4(x,y,z,1), detail/type_mat4x4.inl:536-582: (m0*x + m1*y) + (m2*z + m3*1)
ICX_DI V3 xform(const M4& m, float x, float y, float z) { return (m.c[0] * x + m.c[1] * y) + (m.c[2] * z + m.t); }

// glm::mat3_cast, gtc/quaternion.inl:47-72
ICX_DI M3 mat3_cast(Q4 q) {
    float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z, qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z;
    float qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
    M3 r;
    r.c[0] = v3(1.0f - 2.0f * (qyy + qzz), 2.0f * (qxy + qwz), 2.0f * (qxz - qwy));
    r.c[1] = v3(2.0f * (qxy - qwz), 1.0f - 2.0f * (qxx + qzz), 2.0f * (qyz + qwx));
    r.c[2] = v3(2.0f * (qxz + qwy), 2.0f * (qyz - qwx), 1.0f - 2.0f * (qxx + qyy));
    return r;
}
// quaternion product p*q, detail/type_quat.inl:286-296
ICX_DI Q4 qmul(Q4 p, Q4 q) {
    Q4 r;
    r.w = p.w * q.w - p.x * q.x - p.y * q.y - p.z * q.z;
    r.x = p.w * q.x + p.x * q.w + p.y * q.z - p.z * q.y;
    r.y = p.w * q.y + p.y * q.w + p.z * q.x - p.x * q.z;
    r.z = p.w * q.z + p.z * q.w + p.x * q.y - p.y * q.x;
    return r;
}
// glm::angleAxis, ext/quaternion_trigonometric.inl:30-36.  The reference calls the host libm
// sinf/cosf, which is not correctly rounded and differs between glibc builds (FMA ifuncs);
// here sin/cos are evaluated in double and rounded once, i.e. within 1 ulp of any libm.
ICX_DI Q4 angle_axis(float angle, V3 v) {
    double sd, cd;
    sincos((double)(angle * 0.5f), &sd, &cd);
    float s = (float)sd;
    Q4 r; r.w = (float)cd; r.x = v.x * s; r.y = v.y * s; r.z = v.z * s;
    return r;
}

// ---- full glm 4x4 arithmetic, only for building composite transforms (zero signs matter) ----
struct V4 { float x, y, z, w; };
struct F4 { V4 c[4]; };
ICX_DI V4 v4(float x, float y, float z, float w) { V4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
ICX_DI V4 operator*(V4 a, float s) { return v4(a.x * s, a.y * s, a.z * s, a.w * s); }
ICX_DI V4 operator+(V4 a, V4 b) { return v4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
// glm mat4*mat4, detail/type_mat4x4.inl:630-650 (left-to-right accumulation)
ICX_DI F4 mul(const F4& a, const F4& b) {
    F4 r;
#pragma unroll
    for (int j = 0; j < 4; j++) r.c[j] = a.c[0] * b.c[j].x + a.c[1] * b.c[j].y + a.c[2] * b.c[j].z + a.c[3] * b.c[j].w;
    return r;
}
// TransformComponent::get_model_transform, source/physics/transform_component.h:28-37
ICX_DI F4 model_transform(V3 origin, Q4 q, V3 scale) {
    F4 I; I.c[0] = v4(1, 0, 0, 0); I.c[1] = v4(0, 1, 0, 0); I.c[2] = v4(0, 0, 1, 0); I.c[3] = v4(0, 0, 0, 1);
    F4 tr = I;
    tr.c[3] = I.c[0] * origin.x + I.c[1] * origin.y + I.c[2] * origin.z + I.c[3];   // ext/matrix_transform.inl translate
    M3 r3 = mat3_cast(q);
    F4 rot;
    rot.c[0] = v4(r3.c[0].x, r3.c[0].y, r3.c[0].z, 0); rot.c[1] = v4(r3.c[1].x, r3.c[1].y, r3.c[1].z, 0);
    rot.c[2] = v4(r3.c[2].x, r3.c[2].y, r3.c[2].z, 0); rot.c[3] = v4(0, 0, 0, 1);
    F4 sc;
    sc.c[0] = I.c[0] * scale.x; sc.c[1] = I.c[1] * scale.y; sc.c[2] = I.c[2] * scale.z; sc.c[3] = I.c[3];
    return mul(mul(tr, rot), sc);
}

} // namespace icx
