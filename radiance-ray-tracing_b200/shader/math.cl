// math.cl -- device math library visible to Radiance user programs (B200 backend).
//
// Same names, argument meaning and results as the reference header of the same name
// (radiance/shader/math.cl: random_pcg3d :10-23, MultiplyMat4Vec4 :25-31, MultiplyMat4Mat4 :33-54,
// InverseMat4x4 :56-183, EulerX/Y/ZToMat4x4 :185-252, TransformToTranslate :254-259,
// Vec4ToMat4x4 :261-267, GetNormalSpace :269-298), written from scratch on top of cl_compat.h.
// A mat4x4 is a float16 holding a row-major 4x4: element (r,c) = m.s[4*r + c].
#ifndef RD_B200_MATH_CL
#define RD_B200_MATH_CL

typedef float16 mat4x4;
typedef float3 vec3;
typedef float4 vec4;

// PCG3D hash of Jarzynski & Olano (JCGT 9(3), 2020): the integer stage is exact, the float stage
// divides by (float)0xffffffff like the reference, so results can equal 1.0.
RD_F float3 random_pcg3d(uint3 v)
{
    uint x = v.x * 1664525u + 1013904223u;
    uint y = v.y * 1664525u + 1013904223u;
    uint z = v.z * 1664525u + 1013904223u;
    x += y * z; y += z * x; z += x * y;
    x ^= x >> 16; y ^= y >> 16; z ^= z >> 16;
    x += y * z; y += z * x; z += x * y;
    const float scale = (float)0xffffffffu;
    // correctly rounded on purpose (the reference's CPU path divides exactly; user-program division is otherwise 2.5 ulp)
    return float3(RD_FDIV((float)x, scale), RD_FDIV((float)y, scale), RD_FDIV((float)z, scale));
}

// out = a * b with b a column vector; each row is summed left to right
RD_F void MultiplyMat4Vec4(const mat4x4* a, const vec4* b, vec4* out)
{
    float r[4];
    for (int i = 0; i < 4; ++i)
        r[i] = a->s[4 * i] * b->x + a->s[4 * i + 1] * b->y + a->s[4 * i + 2] * b->z + a->s[4 * i + 3] * b->w;
    out->x = r[0]; out->y = r[1]; out->z = r[2]; out->w = r[3];
}

RD_F void MultiplyMat4Mat4(const mat4x4* a, const mat4x4* b, mat4x4* out)
{
    mat4x4 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            r.s[4 * i + j] = a->s[4 * i] * b->s[j] + a->s[4 * i + 1] * b->s[4 + j] +
                             a->s[4 * i + 2] * b->s[8 + j] + a->s[4 * i + 3] * b->s[12 + j];
    *out = r;
}

// General 4x4 inverse by pairing 2x2 sub-determinants of the upper and lower row pairs.
// Returns false (and leaves *invOut untouched) for a singular matrix.
RD_F bool InverseMat4x4(const mat4x4* m, mat4x4* invOut)
{
    const float* a = m->s;
    const float s0 = a[0] * a[5] - a[4] * a[1], s1 = a[0] * a[6] - a[4] * a[2], s2 = a[0] * a[7] - a[4] * a[3];
    const float s3 = a[1] * a[6] - a[5] * a[2], s4 = a[1] * a[7] - a[5] * a[3], s5 = a[2] * a[7] - a[6] * a[3];
    const float c5 = a[10] * a[15] - a[14] * a[11], c4 = a[9] * a[15] - a[13] * a[11], c3 = a[9] * a[14] - a[13] * a[10];
    const float c2 = a[8] * a[15] - a[12] * a[11], c1 = a[8] * a[14] - a[12] * a[10], c0 = a[8] * a[13] - a[12] * a[9];
    const float det = s0 * c5 - s1 * c4 + s2 * c3 + s3 * c2 - s4 * c1 + s5 * c0;
    if (det == 0.0f) return false;
    const float r = 1.0f / det;
    mat4x4 o;
    o.s[0] = (a[5] * c5 - a[6] * c4 + a[7] * c3) * r;
    o.s[1] = (-a[1] * c5 + a[2] * c4 - a[3] * c3) * r;
    o.s[2] = (a[13] * s5 - a[14] * s4 + a[15] * s3) * r;
    o.s[3] = (-a[9] * s5 + a[10] * s4 - a[11] * s3) * r;
    o.s[4] = (-a[4] * c5 + a[6] * c2 - a[7] * c1) * r;
    o.s[5] = (a[0] * c5 - a[2] * c2 + a[3] * c1) * r;
    o.s[6] = (-a[12] * s5 + a[14] * s2 - a[15] * s1) * r;
    o.s[7] = (a[8] * s5 - a[10] * s2 + a[11] * s1) * r;
    o.s[8] = (a[4] * c4 - a[5] * c2 + a[7] * c0) * r;
    o.s[9] = (-a[0] * c4 + a[1] * c2 - a[3] * c0) * r;
    o.s[10] = (a[12] * s4 - a[13] * s2 + a[15] * s0) * r;
    o.s[11] = (-a[8] * s4 + a[9] * s2 - a[11] * s0) * r;
    o.s[12] = (-a[4] * c3 + a[5] * c1 - a[6] * c0) * r;
    o.s[13] = (a[0] * c3 - a[1] * c1 + a[2] * c0) * r;
    o.s[14] = (-a[12] * s3 + a[13] * s1 - a[14] * s0) * r;
    o.s[15] = (a[8] * s3 - a[9] * s1 + a[10] * s0) * r;
    *invOut = o;
    return true;
}

RD_F void rd_mat_identity(mat4x4* out)
{
    for (int i = 0; i < 16; ++i) out->s[i] = (i % 5 == 0) ? 1.0f : 0.0f;
}

// Rotations about the coordinate axes (angles in radians), column-vector convention.
RD_F void EulerXToMat4x4(float thetaX, mat4x4* out)
{
    const float c = cos(thetaX), s = sin(thetaX);
    rd_mat_identity(out);
    out->s[5] = c; out->s[6] = -s; out->s[9] = s; out->s[10] = c;
}

RD_F void EulerYToMat4x4(float thetaY, mat4x4* out)
{
    const float c = cos(thetaY), s = sin(thetaY);
    rd_mat_identity(out);
    out->s[0] = c; out->s[2] = s; out->s[8] = -s; out->s[10] = c;
}

RD_F void EulerZToMat4x4(float thetaZ, mat4x4* out)
{
    const float c = cos(thetaZ), s = sin(thetaZ);
    rd_mat_identity(out);
    out->s[0] = c; out->s[1] = -s; out->s[4] = s; out->s[5] = c;
}

RD_F void TransformToTranslate(const mat4x4* a, vec3* out)
{
    out->x = a->s[3]; out->y = a->s[7]; out->z = a->s[11];
}

RD_F void Vec4ToMat4x4(vec4 r0, vec4 r1, vec4 r2, vec4 r3, mat4x4* out)
{
    const vec4 rows[4] = {r0, r1, r2, r3};
    for (int i = 0; i < 4; ++i) {
        out->s[4 * i] = rows[i].x; out->s[4 * i + 1] = rows[i].y;
        out->s[4 * i + 2] = rows[i].z; out->s[4 * i + 3] = rows[i].w;
    }
}

// Orthonormal frame with `normal` as its third column: columns = (tangent, bitangent, normal).
// The tangent is normalize(X x normal) unless normal is (anti)parallel to X, then it is Y.
RD_F void GetNormalSpace(float3 normal, mat4x4* out)
{
    const float3 xAxis = float3(1.0f, 0.0f, 0.0f);
    float3 tangent = float3(0.0f, 1.0f, 0.0f);
    if (1.0f - fabs(dot(xAxis, normal)) > 1e-6f) tangent = normalize(cross(xAxis, normal));
    const float3 bitangent = cross(normal, tangent);
    rd_mat_identity(out);
    out->s[0] = tangent.x; out->s[4] = tangent.y; out->s[8] = tangent.z;
    out->s[1] = bitangent.x; out->s[5] = bitangent.y; out->s[9] = bitangent.z;
    out->s[2] = normal.x; out->s[6] = normal.y; out->s[10] = normal.z;
}

#endif
