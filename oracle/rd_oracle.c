/* rd_oracle.c -- plain-C restatement of the reference's traceRay path (see rd_oracle.h).
 * TEST INFRASTRUCTURE ONLY. Compiled with -ffp-contract=off: every operation rounds separately, in
 * the order the reference's expressions are written, so results are bit-identical to the reference
 * sources compiled the same way (oracle/_ref) -- checked by tests/test_oracle_port.py against
 * tests/golden/. */
#include "rd_oracle.h"

#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ---- byte layouts, radiance/src/core.h:34-101 / radiance/shader/data.cl:4-99 ------------------ */
typedef struct { uint32_t type, nodeByteOffset, a, b; } AsHeader; /* top: a=instByteOffset b=totalBufferSize; bottom: a=faceByteOffset b=vertexOffset */
typedef struct { float lo[4], hi[4]; uint32_t u0, u1, u2, u3; } Node; /* inner: u0=left u1=right; leaf: u0=count|0x80000000 u1=start u2=type */
typedef struct { uint32_t i0, i1, i2, primID; } Face;
typedef struct { float m[16]; uint32_t sbt, instanceID, customID, blasOffset; } Inst;
enum { TYPE_INST = 1, TYPE_TRIG = 2, TYPE_TOP_AS = 1, TYPE_BOT_AS = 2 };
#define TOP_STACK 8    /* radiance.cl:38 */
#define TOP_STACK_MAX 256
#define BOT_STACK 100  /* radiance.cl:39 */

/* ---- radiance/shader/math.cl:10-23 ----------------------------------------------------------- */
void rdo_pcg3d(const uint32_t in[3], float out[3])
{
    uint32_t x = in[0] * 1664525u + 1013904223u, y = in[1] * 1664525u + 1013904223u, z = in[2] * 1664525u + 1013904223u;
    x += y * z; y += z * x; z += x * y;
    x ^= x >> 16; y ^= y >> 16; z ^= z >> 16;
    x += y * z; y += z * x; z += x * y;
    out[0] = (float)x / (float)0xffffffffu;
    out[1] = (float)y / (float)0xffffffffu;
    out[2] = (float)z / (float)0xffffffffu;
}

/* ---- radiance/shader/math.cl:56-183: cofactor expansion, six signed triple products per entry,
 * summed left to right. Each row lists the products as sign + three hex element indices, in the
 * order the reference writes them; entry k of the table is inv[ORDER[k]]. --------------------- */
static const char* const COF[16] = {
    "+5af-5be-96f+97e+d6b-d7a", "-4af+4be+86f-87e-c6b+c7a", "+49f-4bd-85f+87d+c5b-c79", "-49e+4ad+85e-86d-c5a+c69",
    "-1af+1be+92f-93e-d2b+d3a", "+0af-0be-82f+83e+c2b-c3a", "-09f+0bd+81f-83d-c1b+c39", "+09e-0ad-81e+82d+c1a-c29",
    "+16f-17e-52f+53e+d27-d36", "-06f+07e+42f-43e-c27+c36", "+05f-07d-41f+43d+c17-c35", "-05e+06d+41e-42d-c16+c25",
    "-16b+17a+52b-53a-927+936", "+06b-07a-42b+43a+827-836", "-05b+079+41b-439-817+835", "+05a-069-41a+429+816-825"};
static const int ORDER[16] = {0, 4, 8, 12, 1, 5, 9, 13, 2, 6, 10, 14, 3, 7, 11, 15};

static int hexv(char c) { return c <= '9' ? c - '0' : c - 'a' + 10; }

int rdo_inverse4x4(const float m[16], float out[16])
{
    float inv[16];
    for (int k = 0; k < 16; ++k) {
        const char* s = COF[k];
        float acc = 0.0f;
        for (int t = 0; t < 6; ++t, s += 4) {
            float a = m[hexv(s[1])];
            if (t == 0 && s[0] == '-') a = -a;               /* leading unary minus binds to the first factor */
            const float prod = a * m[hexv(s[2])] * m[hexv(s[3])];
            if (t == 0) acc = prod;
            else if (s[0] == '+') acc = acc + prod;
            else acc = acc - prod;
        }
        inv[ORDER[k]] = acc;
    }
    float det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    if (det == 0) return 0;
    det = 1.0f / det;
    for (int i = 0; i < 16; ++i) out[i] = inv[i] * det;
    return 1;
}

/* ---- radiance/shader/radiance.cl:195-208 ------------------------------------------------------ */
static float fmin2(float a, float b) { return b < a ? b : a; }   /* OpenCL min/max as in the compat prelude */
static float fmax2(float a, float b) { return a < b ? b : a; }

int rdo_intersect_aabb(const float o[3], const float d[3], const float lo[3], const float hi[3])
{
    float t1[3], t2[3];
    for (int a = 0; a < 3; ++a) {
        const float tmin = (lo[a] - o[a]) / d[a], tmax = (hi[a] - o[a]) / d[a];
        t1[a] = fmin2(tmin, tmax);
        t2[a] = fmax2(tmin, tmax);
    }
    const float tNear = fmax2(fmax2(t1[0], t1[1]), t1[2]);
    const float tFar = fmin2(fmin2(t2[0], t2[1]), t2[2]);
    return tFar > fmax2(tNear, 0.0f);
}

/* ---- radiance/shader/radiance.cl:211-251 ------------------------------------------------------ */
static void cross3(const float a[3], const float b[3], float r[3])
{
    r[0] = a[1] * b[2] - a[2] * b[1];
    r[1] = a[2] * b[0] - a[0] * b[2];
    r[2] = a[0] * b[1] - a[1] * b[0];
}
static float dot3(const float a[3], const float b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

int rdo_intersect_triangle(const float o[3], const float d[3], const float v[9], float out[7])
{
    float e1[3], e2[3], p[3], s[3], q[3];
    for (int a = 0; a < 3; ++a) { e1[a] = v[3 + a] - v[a]; e2[a] = v[6 + a] - v[a]; }
    cross3(d, e2, p);
    const float det = dot3(e1, p);
    if (det == 0) return 0;
    const float inv = 1.0f / det;
    for (int a = 0; a < 3; ++a) s[a] = o[a] - v[a];
    const float b1 = inv * dot3(s, p);
    cross3(s, e1, q);
    const float b2 = inv * dot3(d, q);
    const float t = inv * dot3(e2, q);
    if (b1 < 0 || b1 > 1) return 0;
    if (b2 < 0 || b1 + b2 > 1) return 0;
    if (!(t > 0)) return 0;
    out[0] = t;
    for (int a = 0; a < 3; ++a) out[1 + a] = o[a] + d[a] * t;
    out[4] = 1 - b1 - b2; out[5] = b1; out[6] = b2;
    return 1;
}

/* ---- radiance/shader/radiance.cl:41-108 (intersectBot) ------------------------------------------ */
typedef struct { rdo_hit* hit; float transform[16]; int terminate; int cont; int topStack; int overflow; } TraceState;

static int intersect_bot(const char* as, const float o[3], const float d[3], float Tmin, float Tmax, TraceState* st)
{
    const AsHeader* hdr = (const AsHeader*)as;
    const Node* nodes = (const Node*)(as + hdr->nodeByteOffset);
    const Face* faces = (const Face*)(as + hdr->a);
    const float* verts = (const float*)(as + hdr->b);   /* float4 per vertex */
    int hasIntersected = 0;
    uint32_t stack[BOT_STACK + 2];
    int sp = 0;
    stack[sp++] = 0;
    while (sp) {
        const Node* node = nodes + stack[--sp];
        if (!(node->u0 & 0x80000000u)) {
            if (rdo_intersect_aabb(o, d, node->lo, node->hi)) {
                stack[sp++] = node->u1;   /* right first, so the left child is popped first */
                stack[sp++] = node->u0;
                if (sp > BOT_STACK) return 0;   /* the reference prints and reports a miss */
            }
        } else if (node->u2 == TYPE_TRIG) {
            const uint32_t count = node->u0 & 0x7fffffffu;
            for (uint32_t i = 0; i < count; ++i) {
                const Face* f = faces + node->u1 + i;
                float v[9], r[7];
                for (int a = 0; a < 3; ++a) { v[a] = verts[4 * f->i0 + a]; v[3 + a] = verts[4 * f->i1 + a]; v[6 + a] = verts[4 * f->i2 + a]; }
                if (rdo_intersect_triangle(o, d, v, r) && r[0] < st->hit->t && r[0] > Tmin && r[0] < Tmax) {
                    st->hit->t = r[0];
                    memcpy(st->hit->point, r + 1, 12);
                    memcpy(st->hit->bary, r + 4, 12);
                    st->hit->prim = f->primID;
                    hasIntersected = 1;
                    st->hit->anyhit_calls++;              /* callAnyHit, radiance.cl:99 */
                    if (st->terminate) st->cont = 0;
                    if (!st->cont) return hasIntersected;
                }
            }
        }
    }
    return hasIntersected;
}

static void mat_vec(const float m[16], const float b[4], float out[4])   /* math.cl:25-31 */
{
    for (int r = 0; r < 4; ++r) out[r] = m[4 * r] * b[0] + m[4 * r + 1] * b[1] + m[4 * r + 2] * b[2] + m[4 * r + 3] * b[3];
}

/* ---- radiance/shader/radiance.cl:110-192 (intersectTop) ----------------------------------------- */
static int intersect_top(const char* as, const float o[3], const float d[3], float Tmin, float Tmax, TraceState* st)
{
    const AsHeader* hdr = (const AsHeader*)as;
    const Node* nodes = (const Node*)(as + hdr->nodeByteOffset);
    const Inst* insts = (const Inst*)(as + hdr->a);
    int hasIntersected = 0;
    uint32_t stack[TOP_STACK_MAX + 2];
    int sp = 0;
    stack[sp++] = 0;
    st->cont = 1;
    while (sp) {
        const Node* node = nodes + stack[--sp];
        if (!(node->u0 & 0x80000000u)) {
            if (rdo_intersect_aabb(o, d, node->lo, node->hi)) {
                stack[sp++] = node->u1;
                stack[sp++] = node->u0;
                if (sp > st->topStack) { st->overflow = 1; return 0; }   /* the reference prints and reports a miss */
            }
        } else if (node->u2 == TYPE_INST) {
            const uint32_t count = node->u0 & 0x7fffffffu;
            for (uint32_t i = 0; i < count; ++i) {
                const Inst* in = insts + node->u1 + i;
                rdo_hit saved = *st->hit;
                float inv[16], pos[4] = {o[0], o[1], o[2], 1.0f}, dir[4] = {d[0], d[1], d[2], 0.0f}, lo[4], ld[4];
                if (!rdo_inverse4x4(in->m, inv)) memset(inv, 0, sizeof(inv));   /* the reference leaves `inverse` unset here */
                mat_vec(inv, pos, lo);
                mat_vec(inv, dir, ld);
                st->hit->inst = in->instanceID; st->hit->custom = in->customID; st->hit->sbt = in->sbt;
                const int result = intersect_bot(as + in->blasOffset, lo, ld, Tmin, Tmax, st);
                hasIntersected = hasIntersected || result;
                if (!st->cont) return hasIntersected;
                if (!result) { st->hit->inst = saved.inst; st->hit->custom = saved.custom; st->hit->sbt = saved.sbt; }
            }
        }
    }
    return hasIntersected;
}

typedef struct { const char* blob; const float* rays; size_t b, e; int terminate; rdo_hit* out; int topStack; unsigned char* overflow; } Job;

static void* trace_job(void* arg)
{
    Job* j = (Job*)arg;
    for (size_t i = j->b; i < j->e; ++i) {
        const float* r = j->rays + 8 * i;
        rdo_hit h;
        memset(&h, 0, sizeof(h));
        h.t = FLT_MAX;                                          /* radiance.cl:265 */
        TraceState st;
        st.hit = &h; st.terminate = j->terminate; st.cont = 1; st.topStack = j->topStack; st.overflow = 0;
        const float o[3] = {r[0], r[1], r[2]}, d[3] = {r[4], r[5], r[6]};
        h.hit = (uint32_t)intersect_top(j->blob, o, d, r[3], r[7], &st);
        if (!h.hit) { const uint32_t calls = h.anyhit_calls; memset(&h, 0, sizeof(h)); h.anyhit_calls = calls; }
        j->out[i] = h;
        if (j->overflow) j->overflow[i] = (unsigned char)st.overflow;
    }
    return 0;
}

/* top_stack: entries of the top-level stack. TOP_STACK (8) is the reference (radiance.cl:38), whose walk gives up -- and reports a
 * miss -- when an instance hierarchy needs more (4096 instances do); a larger value walks the same blob in the same order without
 * that limit, and overflow[i] (optional) tells which rays the limit would have cut. */
void rdo_trace_ex(const void* tlas_blob, const float* rays, size_t n, int terminate_on_first_hit, rdo_hit* out, int nthreads,
                  int top_stack, unsigned char* overflow);
void rdo_trace(const void* tlas_blob, const float* rays, size_t n, int terminate_on_first_hit, rdo_hit* out, int nthreads)
{
    rdo_trace_ex(tlas_blob, rays, n, terminate_on_first_hit, out, nthreads, TOP_STACK, 0);
}

void rdo_trace_ex(const void* tlas_blob, const float* rays, size_t n, int terminate_on_first_hit, rdo_hit* out, int nthreads,
                  int top_stack, unsigned char* overflow)
{
    if (top_stack < 1) top_stack = 1;
    if (top_stack > TOP_STACK_MAX) top_stack = TOP_STACK_MAX;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    if (n < 256) nthreads = 1;
    pthread_t th[256];
    Job jobs[256];
    const size_t chunk = (n + (size_t)nthreads - 1) / (size_t)nthreads;
    int started = 0;
    for (int t = 0; t < nthreads; ++t) {
        Job* j = &jobs[t];
        j->blob = (const char*)tlas_blob; j->rays = rays; j->terminate = terminate_on_first_hit; j->out = out;
        j->topStack = top_stack; j->overflow = overflow;
        j->b = (size_t)t * chunk; j->e = j->b + chunk < n ? j->b + chunk : n;
        if (j->b >= j->e) break;
        if (nthreads == 1) { trace_job(j); return; }
        pthread_create(&th[t], 0, trace_job, j);
        started++;
    }
    for (int t = 0; t < started; ++t) pthread_join(th[t], 0);
}

/* ---- blob builders (layout of radiance/src/radiance.cpp:318-425, data.cl:237-278) ---------------- */
typedef struct { float lo[3], hi[3], c[3]; uint32_t id; } Prim;
typedef struct { Node* nodes; uint32_t count, cap; uint32_t* order; uint32_t norder; int leafType; } Builder;

static int g_axis;
static int cmp_prim(const void* a, const void* b)
{
    const float x = ((const Prim*)a)->c[g_axis], y = ((const Prim*)b)->c[g_axis];
    return x < y ? -1 : (x > y ? 1 : 0);
}

static uint32_t build_rec(Builder* B, Prim* p, uint32_t n)
{
    if (B->count == B->cap) { B->cap = B->cap ? B->cap * 2 : 64; B->nodes = (Node*)realloc(B->nodes, B->cap * sizeof(Node)); }
    const uint32_t me = B->count++;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = 0; i < n; ++i)
        for (int a = 0; a < 3; ++a) { if (p[i].lo[a] < lo[a]) lo[a] = p[i].lo[a]; if (p[i].hi[a] > hi[a]) hi[a] = p[i].hi[a]; }
    Node nd;
    memset(&nd, 0, sizeof(nd));
    memcpy(nd.lo, lo, 12); memcpy(nd.hi, hi, 12);
    if (n < 8) {                                            /* MAX_LEAF_PRIM_SIZE, radiance/src/bvh.h:9 */
        nd.u0 = n | 0x80000000u; nd.u1 = B->norder; nd.u2 = (uint32_t)B->leafType;
        for (uint32_t i = 0; i < n; ++i) B->order[B->norder++] = p[i].id;
        B->nodes[me] = nd;
        return me;
    }
    int axis = 0;
    for (int a = 1; a < 3; ++a) if (hi[a] - lo[a] > hi[axis] - lo[axis]) axis = a;
    g_axis = axis;
    qsort(p, n, sizeof(Prim), cmp_prim);
    const uint32_t half = n / 2;
    const uint32_t l = build_rec(B, p, half);
    const uint32_t r = build_rec(B, p + half, n - half);
    nd.u0 = l; nd.u1 = r;
    B->nodes[me] = nd;
    return me;
}

void* rdo_blas_build(const float* verts, uint32_t nv, const uint32_t* idx, uint32_t nt, size_t* size)
{
    Prim* p = (Prim*)malloc((nt ? nt : 1) * sizeof(Prim));
    for (uint32_t i = 0; i < nt; ++i) {
        for (int a = 0; a < 3; ++a) {
            const float x = verts[3 * idx[3 * i] + a], y = verts[3 * idx[3 * i + 1] + a], z = verts[3 * idx[3 * i + 2] + a];
            p[i].lo[a] = fminf(fminf(x, y), z); p[i].hi[a] = fmaxf(fmaxf(x, y), z);
            p[i].c[a] = (x + y + z) / 3.0f;
        }
        p[i].id = i;
    }
    Builder B;
    memset(&B, 0, sizeof(B));
    B.order = (uint32_t*)malloc((nt ? nt : 1) * sizeof(uint32_t));
    B.leafType = TYPE_TRIG;
    build_rec(&B, p, nt);
    const size_t nodeBytes = (size_t)B.count * sizeof(Node), faceBytes = (size_t)nt * sizeof(Face), vertBytes = (size_t)nv * 16;
    const size_t total = sizeof(AsHeader) + nodeBytes + faceBytes + vertBytes;
    char* blob = (char*)calloc(total, 1);
    AsHeader* h = (AsHeader*)blob;
    h->type = TYPE_BOT_AS; h->nodeByteOffset = sizeof(AsHeader);
    h->a = (uint32_t)(sizeof(AsHeader) + nodeBytes); h->b = (uint32_t)(sizeof(AsHeader) + nodeBytes + faceBytes);
    memcpy(blob + h->nodeByteOffset, B.nodes, nodeBytes);
    Face* f = (Face*)(blob + h->a);
    for (uint32_t k = 0; k < nt; ++k) {
        const uint32_t t = B.order[k];
        f[k].i0 = idx[3 * t]; f[k].i1 = idx[3 * t + 1]; f[k].i2 = idx[3 * t + 2]; f[k].primID = t;
    }
    float* dv = (float*)(blob + h->b);
    for (uint32_t i = 0; i < nv; ++i) { dv[4 * i] = verts[3 * i]; dv[4 * i + 1] = verts[3 * i + 1]; dv[4 * i + 2] = verts[3 * i + 2]; }
    free(p); free(B.order); free(B.nodes);
    if (size) *size = total;
    return blob;
}

void* rdo_tlas_build(const rdo_instance* inst, uint32_t n, size_t* size)
{
    Prim* p = (Prim*)malloc((n ? n : 1) * sizeof(Prim));
    for (uint32_t i = 0; i < n; ++i) {
        /* world box = transform of the 8 corners of the bottom-level root box (bvh.cpp:370-401) */
        const AsHeader* bh = (const AsHeader*)inst[i].blas;
        const Node* root = (const Node*)((const char*)inst[i].blas + bh->nodeByteOffset);
        for (int a = 0; a < 3; ++a) { p[i].lo[a] = FLT_MAX; p[i].hi[a] = -FLT_MAX; }
        for (int c = 0; c < 8; ++c) {
            const float v[4] = {(c & 1) ? root->hi[0] : root->lo[0], (c & 2) ? root->hi[1] : root->lo[1], (c & 4) ? root->hi[2] : root->lo[2], 1.0f};
            float w[4];
            mat_vec(inst[i].transform, v, w);
            for (int a = 0; a < 3; ++a) { if (w[a] < p[i].lo[a]) p[i].lo[a] = w[a]; if (w[a] > p[i].hi[a]) p[i].hi[a] = w[a]; }
        }
        for (int a = 0; a < 3; ++a) p[i].c[a] = 0.5f * (p[i].lo[a] + p[i].hi[a]);
        p[i].id = i;
    }
    Builder B;
    memset(&B, 0, sizeof(B));
    B.order = (uint32_t*)malloc((n ? n : 1) * sizeof(uint32_t));
    B.leafType = TYPE_INST;
    build_rec(&B, p, n);
    /* unique bottom-level blobs in first-seen order, appended after [header | nodes | instances] */
    const size_t nodeBytes = (size_t)B.count * sizeof(Node), instBytes = (size_t)n * sizeof(Inst);
    size_t total = sizeof(AsHeader) + nodeBytes + instBytes;
    uint32_t* off = (uint32_t*)malloc((n ? n : 1) * sizeof(uint32_t));
    for (uint32_t k = 0; k < n; ++k) {
        const uint32_t i = B.order[k];
        uint32_t found = 0xffffffffu;
        for (uint32_t m = 0; m < k; ++m) if (inst[B.order[m]].blas == inst[i].blas) { found = off[m]; break; }
        if (found == 0xffffffffu) { found = (uint32_t)total; total += inst[i].blas_size; }
        off[k] = found;
    }
    char* blob = (char*)calloc(total, 1);
    AsHeader* h = (AsHeader*)blob;
    h->type = TYPE_TOP_AS; h->nodeByteOffset = sizeof(AsHeader);
    h->a = (uint32_t)(sizeof(AsHeader) + nodeBytes); h->b = (uint32_t)total;
    memcpy(blob + h->nodeByteOffset, B.nodes, nodeBytes);
    Inst* di = (Inst*)(blob + h->a);
    for (uint32_t k = 0; k < n; ++k) {
        const uint32_t i = B.order[k];
        memcpy(di[k].m, inst[i].transform, 64);
        di[k].sbt = inst[i].sbt; di[k].instanceID = i; di[k].customID = inst[i].custom; di[k].blasOffset = off[k];
        memcpy(blob + off[k], inst[i].blas, inst[i].blas_size);
    }
    free(p); free(B.order); free(B.nodes); free(off);
    if (size) *size = total;
    return blob;
}

void rdo_free(void* p) { free(p); }
