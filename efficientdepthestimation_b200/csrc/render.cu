// render.cu -- depth-mesh rendering as a CUDA rasteriser: no graphics API.
//
// replaces (reference, executed through OpenGL / GLSL 4.60 in a GLFW window):
//   mesh from depth      DepthRenderer/DepthRenderer/render.py:465-546 (+ Benchmark/benchmark.py:949 displacement)
//   vertex shader        DepthRenderer/DepthRenderer/shaders/shader.vert:1-14   gl_Position = mvp * vec4(p, 1)
//   rasterisation        fixed function: back-face culling, GL_LESS depth test, clear to opaque black (render.py:631-634,807)
//   fragment shader      shaders/shader.frag:1-8   texture2D(colourSampler, uv), GL_LINEAR + GL_CLAMP (render.py:357-360)
//   read-back            glReadPixels RGBA8, row 0 = bottom of the window (render.py:787)
//
// Rules the GL specification leaves open are fixed as in oracle/render_cpu.py (the CPU restatement this file matches bit
// for bit): window coordinates snapped to 1/256 pixel, integer edge functions with the top-left fill rule, fp32 window
// depth with ties going to the triangle drawn first, perspective-correct uv, clamp-to-edge bilinear fetch, round-to-
// nearest-even colour.  Every fp32 / fp64 operation below is an explicitly rounded intrinsic: no FMA contraction.
//
// Structure (HBM / atomics bound integer work, no tensor cores): many frames of ONE mesh per call (the benchmark renders
// 300 camera poses per depth map), grid = frames x work items:
//   vertex_kernel      [frame][vertex] -> snapped window position, depth, 1/w                     (16 B per vertex)
//   raster_kernel      one thread per (frame, triangle): walks the bounding box, atomicMin of (depth bits << 32 | triangle)
//                      into the 64-bit z/id buffer; triangles with large boxes go to a list ...
//   raster_big_kernel  ... that whole blocks walk cooperatively (coarse meshes, density 0..4)
//   shade_kernel       one thread per (frame, pixel): re-derives the winner's barycentrics, fetches the texture, writes RGBA8
#include <algorithm>

#include "common.cuh"

namespace ede {

constexpr int R_SUB = 256;                       // sub-pixel grid
constexpr int R_HALF = R_SUB / 2;
constexpr int R_BIG_AREA = 1024;                 // bounding boxes above this many pixels are walked by a whole block
constexpr unsigned long long R_EMPTY = ~0ull;

struct RVertex { int xi, yi; float zw, inv_w; };  // xi == INT_MIN: clip w <= 0 (triangle dropped, no near clipping)

// ------------------------------------------------------------------ mesh from depth (render.py:465-546)
__global__ void mesh_from_depth_kernel(const EdeMeshFromDepth a) {
    const int n = (1 << a.density) + 1;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * n) return;
    const int row = i / n, col = i - row * n;
    const double step = 2.0 / (double)(n - 1);                       // np.linspace(-1, 1, n): start + i * step in fp64
    const float x = col == n - 1 ? 1.f : (float)__dadd_rn(-1.0, __dmul_rn((double)col, step));
    const float y0 = row == n - 1 ? -1.f : (float)__dadd_rn(1.0, __dmul_rn((double)row, -step));
    const double r = __ddiv_rn((double)a.h, (double)a.w);
    // y = (height / width) * y - 0.5 * (1.0 - height / width) * y: the Python scalars are evaluated in fp64, then take
    // the float32 dtype of the linspace array, so the two products and the difference are fp32 operations
    const float yy = __fsub_rn(__fmul_rn((float)r, y0), __fmul_rn((float)__dmul_rn(0.5, __dsub_rn(1.0, r)), y0));
    const double tstep = 1.0 / (double)(n - 1);
    const float ut = col == n - 1 ? 1.f : (float)__dmul_rn((double)col, tstep);
    const float vt = row == n - 1 ? 0.f : (float)__dadd_rn(1.0, __dmul_rn((double)row, -tstep));
    // nearest depth pixel: u = int(col / n * width), v = int((1 - row / n) * height - 1)
    const long long u = (long long)__dmul_rn(__ddiv_rn((double)col, (double)n), (double)a.w);
    const long long v = (long long)__dsub_rn(__dmul_rn(__dsub_rn(1.0, __ddiv_rn((double)row, (double)n)), (double)a.h), 1.0);
    const int vv = (int)(v < 0 ? v + a.h : v);                       // numpy wraps a negative index (never below -1 here)
    const double d = (double)a.depth[((long long)vv * a.w + u) * a.depth_stride];
    const float z = (float)__dsub_rn(1.0, __ddiv_rn(d, 255.0));
    a.vertices[3 * i] = x;
    a.vertices[3 * i + 1] = yy;
    a.vertices[3 * i + 2] = __fmul_rn(z, a.displacement);            // vertices[:, 2] *= displacement (benchmark.py:949)
    a.uv[2 * i] = ut;
    a.uv[2 * i + 1] = vt;
    if (row < n - 1 && col < n - 1) {                                // triangles (a, b, c), (c, b, d)
        const int q = row * (n - 1) + col;
        const unsigned va = row * n + col, vb = (row + 1) * n + col;
        unsigned* o = a.indices + 6 * (long long)q;
        o[0] = va; o[1] = vb; o[2] = va + 1; o[3] = va + 1; o[4] = vb; o[5] = vb + 1;
    }
}

// ------------------------------------------------------------------ vertex stage (shader.vert:11 + viewport)
__device__ __forceinline__ float mvp_row(const float* m, float x, float y, float z) {
    return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[0], x), __fmul_rn(m[1], y)), __fmul_rn(m[2], z)), m[3]);
}
__device__ __forceinline__ int snap(float w) {
    const float s = fminf(fmaxf(__fmul_rn(w, (float)R_SUB), -16777216.f), 16777216.f);
    return __float2int_rn(s);                                        // NaN -> 0 (only on dropped vertices)
}

__global__ void render_vertex_kernel(const EdeRender a, RVertex* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int f = blockIdx.y;
    if (i >= a.n_vertices) return;
    const float* m = a.mvp + 16 * (long long)f;
    const float x = a.vertices[3 * i], y = a.vertices[3 * i + 1], z = a.vertices[3 * i + 2];
    const float cx = mvp_row(m, x, y, z), cy = mvp_row(m + 4, x, y, z), cz = mvp_row(m + 8, x, y, z),
                cw = mvp_row(m + 12, x, y, z);
    RVertex v;
    v.inv_w = __fdiv_rn(1.f, cw);
    const float xw = __fmul_rn(__fadd_rn(__fdiv_rn(cx, cw), 1.f), 0.5f * (float)a.width);
    const float yw = __fmul_rn(__fadd_rn(__fdiv_rn(cy, cw), 1.f), 0.5f * (float)a.height);
    v.zw = __fadd_rn(__fmul_rn(__fdiv_rn(cz, cw), 0.5f), 0.5f);
    v.xi = cw > 0.f ? snap(xw) : INT_MIN;
    v.yi = snap(yw);
    out[(long long)f * a.n_vertices + i] = v;
}

__global__ void render_clear_kernel(unsigned long long* key, long long n, int* big_count) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        key[i] = R_EMPTY;
    if (blockIdx.x == 0 && threadIdx.x == 0) *big_count = 0;
}

// ------------------------------------------------------------------ rasterisation
struct TriSetup {
    long long x0, y0, x1, y1, x2, y2, area;
    int b0, b1, b2;                                                   // top-left biases of the edges opposite v0, v1, v2
    int px_lo, px_hi, py_lo, py_hi;
    float z0, z1, z2;
};

__device__ __forceinline__ int edge_bias(long long ax, long long ay, long long bx, long long by) {
    const long long dx = bx - ax, dy = by - ay;
    return (dy < 0 || (dy == 0 && dx < 0)) ? 0 : -1;
}
__device__ __forceinline__ long long floor_div(long long a, long long b) {       // b > 0
    const long long q = a / b;
    return (a % b != 0 && a < 0) ? q - 1 : q;
}

__device__ __forceinline__ bool tri_setup(const RVertex& v0, const RVertex& v1, const RVertex& v2, int width, int height,
                                          TriSetup& t) {
    if (v0.xi == INT_MIN || v1.xi == INT_MIN || v2.xi == INT_MIN) return false;
    t.x0 = v0.xi; t.y0 = v0.yi; t.x1 = v1.xi; t.y1 = v1.yi; t.x2 = v2.xi; t.y2 = v2.yi;
    t.area = (t.x1 - t.x0) * (t.y2 - t.y0) - (t.y1 - t.y0) * (t.x2 - t.x0);
    if (t.area <= 0) return false;                                    // GL_CULL_FACE (back faces) + degenerate
    const long long xmin = min(t.x0, min(t.x1, t.x2)), xmax = max(t.x0, max(t.x1, t.x2));
    const long long ymin = min(t.y0, min(t.y1, t.y2)), ymax = max(t.y0, max(t.y1, t.y2));
    t.px_lo = (int)max(0ll, -floor_div(-(xmin - R_HALF), R_SUB));     // ceil((min - 128) / 256)
    t.px_hi = (int)min((long long)width - 1, floor_div(xmax - R_HALF, R_SUB));
    t.py_lo = (int)max(0ll, -floor_div(-(ymin - R_HALF), R_SUB));
    t.py_hi = (int)min((long long)height - 1, floor_div(ymax - R_HALF, R_SUB));
    if (t.px_lo > t.px_hi || t.py_lo > t.py_hi) return false;
    t.b0 = edge_bias(t.x1, t.y1, t.x2, t.y2);
    t.b1 = edge_bias(t.x2, t.y2, t.x0, t.y0);
    t.b2 = edge_bias(t.x0, t.y0, t.x1, t.y1);
    t.z0 = v0.zw; t.z1 = v1.zw; t.z2 = v2.zw;
    return true;
}

__device__ __forceinline__ void tri_pixel(const TriSetup& t, int px, int py, unsigned tri, unsigned long long* key_row0,
                                          int width) {
    const long long cx = (long long)px * R_SUB + R_HALF, cy = (long long)py * R_SUB + R_HALF;
    const long long w0 = (t.x2 - t.x1) * (cy - t.y1) - (t.y2 - t.y1) * (cx - t.x1);
    const long long w1 = (t.x0 - t.x2) * (cy - t.y2) - (t.y0 - t.y2) * (cx - t.x2);
    const long long w2 = (t.x1 - t.x0) * (cy - t.y0) - (t.y1 - t.y0) * (cx - t.x0);
    if (w0 + t.b0 < 0 || w1 + t.b1 < 0 || w2 + t.b2 < 0) return;
    const float a = __ll2float_rn(t.area);
    const float b0 = __fdiv_rn(__ll2float_rn(w0), a), b1 = __fdiv_rn(__ll2float_rn(w1), a), b2 = __fdiv_rn(__ll2float_rn(w2), a);
    const float z = __fadd_rn(__fadd_rn(__fmul_rn(b0, t.z0), __fmul_rn(b1, t.z1)), __fmul_rn(b2, t.z2));
    if (!(z >= 0.f && z <= 1.f)) return;                              // outside the depth range: clipped
    const unsigned long long k = ((unsigned long long)__float_as_uint(z) << 32) | tri;
    atomicMin(key_row0 + (long long)py * width + px, k);              // GL_LESS; equal depth: the earlier triangle wins
}

__global__ void __launch_bounds__(256) render_raster_kernel(const EdeRender a, const RVertex* vtx, unsigned long long* key,
                                                           int* big_count, int2* big_list) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int f = blockIdx.y;
    if (t >= a.n_triangles) return;
    const RVertex* v = vtx + (long long)f * a.n_vertices;
    const unsigned i0 = a.indices[3 * t], i1 = a.indices[3 * t + 1], i2 = a.indices[3 * t + 2];
    TriSetup s;
    if (!tri_setup(v[i0], v[i1], v[i2], a.width, a.height, s)) return;
    const long long box = (long long)(s.px_hi - s.px_lo + 1) * (s.py_hi - s.py_lo + 1);
    if (box > R_BIG_AREA) {
        big_list[atomicAdd(big_count, 1)] = make_int2(f, t);
        return;
    }
    unsigned long long* kf = key + (long long)f * a.width * a.height;
    for (int py = s.py_lo; py <= s.py_hi; ++py)
        for (int px = s.px_lo; px <= s.px_hi; ++px) tri_pixel(s, px, py, (unsigned)t, kf, a.width);
}

__global__ void __launch_bounds__(256) render_raster_big_kernel(const EdeRender a, const RVertex* vtx, unsigned long long* key,
                                                               const int* big_count, const int2* big_list) {
    const int n = *big_count;
    for (int e = blockIdx.x; e < n; e += gridDim.x) {
        const int f = big_list[e].x, t = big_list[e].y;
        const RVertex* v = vtx + (long long)f * a.n_vertices;
        TriSetup s;
        if (!tri_setup(v[a.indices[3 * t]], v[a.indices[3 * t + 1]], v[a.indices[3 * t + 2]], a.width, a.height, s)) continue;
        unsigned long long* kf = key + (long long)f * a.width * a.height;
        const int bw = s.px_hi - s.px_lo + 1;
        const long long box = (long long)bw * (s.py_hi - s.py_lo + 1);
        for (long long i = threadIdx.x; i < box; i += blockDim.x)
            tri_pixel(s, s.px_lo + (int)(i % bw), s.py_lo + (int)(i / bw), (unsigned)t, kf, a.width);
    }
}

// ------------------------------------------------------------------ shading (shader.frag:7, GL_LINEAR + clamp to edge)
__device__ __forceinline__ float lerp_rn(float p, float q, float one_minus_t, float t) {
    return __fadd_rn(__fmul_rn(p, one_minus_t), __fmul_rn(q, t));
}

__global__ void __launch_bounds__(256) render_shade_kernel(const EdeRender a, const RVertex* vtx, const unsigned long long* key) {
    const long long npx = (long long)a.width * a.height;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int f = blockIdx.y;
    if (i >= npx) return;
    const int py = (int)(i / a.width), px = (int)(i - (long long)py * a.width);
    const unsigned long long k = key[(long long)f * npx + i];
    const int out_row = a.flip_y ? a.height - 1 - py : py;
    uchar4* dst = reinterpret_cast<uchar4*>(a.frames) + ((long long)f * a.height + out_row) * a.width + px;
    if (k == R_EMPTY) { *dst = make_uchar4(0, 0, 0, 255); return; }  // glClearColor(0, 0, 0, 1)
    const unsigned t = (unsigned)(k & 0xffffffffu);
    const unsigned i0 = a.indices[3 * t], i1 = a.indices[3 * t + 1], i2 = a.indices[3 * t + 2];
    const RVertex* v = vtx + (long long)f * a.n_vertices;
    const RVertex v0 = v[i0], v1 = v[i1], v2 = v[i2];
    const long long x0 = v0.xi, y0 = v0.yi, x1 = v1.xi, y1 = v1.yi, x2 = v2.xi, y2 = v2.yi;
    const long long cx = (long long)px * R_SUB + R_HALF, cy = (long long)py * R_SUB + R_HALF;
    const float area = __ll2float_rn((x1 - x0) * (y2 - y0) - (y1 - y0) * (x2 - x0));
    const float q0 = __fmul_rn(__fdiv_rn(__ll2float_rn((x2 - x1) * (cy - y1) - (y2 - y1) * (cx - x1)), area), v0.inv_w);
    const float q1 = __fmul_rn(__fdiv_rn(__ll2float_rn((x0 - x2) * (cy - y2) - (y0 - y2) * (cx - x2)), area), v1.inv_w);
    const float q2 = __fmul_rn(__fdiv_rn(__ll2float_rn((x1 - x0) * (cy - y0) - (y1 - y0) * (cx - x0)), area), v2.inv_w);
    const float qs = __fadd_rn(__fadd_rn(q0, q1), q2);
    const float* uv = a.uv;
    const float tu = __fdiv_rn(__fadd_rn(__fadd_rn(__fmul_rn(q0, uv[2 * i0]), __fmul_rn(q1, uv[2 * i1])), __fmul_rn(q2, uv[2 * i2])), qs);
    const float tv = __fdiv_rn(__fadd_rn(__fadd_rn(__fmul_rn(q0, uv[2 * i0 + 1]), __fmul_rn(q1, uv[2 * i1 + 1])), __fmul_rn(q2, uv[2 * i2 + 1])), qs);
    const float fx = __fsub_rn(__fmul_rn(tu, (float)a.tex_w), 0.5f), fy = __fsub_rn(__fmul_rn(tv, (float)a.tex_h), 0.5f);
    const float x0f = floorf(fx), y0f = floorf(fy);
    const float ax = __fsub_rn(fx, x0f), ay = __fsub_rn(fy, y0f);
    const int tx = (int)fminf(fmaxf(x0f, -1.f), (float)a.tex_w), ty = (int)fminf(fmaxf(y0f, -1.f), (float)a.tex_h);
    const int xa = min(max(tx, 0), a.tex_w - 1), xb = min(max(tx + 1, 0), a.tex_w - 1);
    const int ya = min(max(ty, 0), a.tex_h - 1), yb = min(max(ty + 1, 0), a.tex_h - 1);
    const uchar4* tex = reinterpret_cast<const uchar4*>(a.texture);
    const uchar4 c00 = tex[(long long)ya * a.tex_w + xa], c10 = tex[(long long)ya * a.tex_w + xb];
    const uchar4 c01 = tex[(long long)yb * a.tex_w + xa], c11 = tex[(long long)yb * a.tex_w + xb];
    const float omx = __fsub_rn(1.f, ax), omy = __fsub_rn(1.f, ay);
    auto chan = [&](unsigned char p00, unsigned char p10, unsigned char p01, unsigned char p11) {
        const float top = lerp_rn((float)p00, (float)p10, omx, ax), bot = lerp_rn((float)p01, (float)p11, omx, ax);
        const float c = lerp_rn(top, bot, omy, ay);
        return (unsigned char)min(max(__float2int_rn(c), 0), 255);
    };
    *dst = make_uchar4(chan(c00.x, c10.x, c01.x, c11.x), chan(c00.y, c10.y, c01.y, c11.y), chan(c00.z, c10.z, c01.z, c11.z),
                       chan(c00.w, c10.w, c01.w, c11.w));
}

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace ede

extern "C" {

int ede_mesh_from_depth(const EdeMeshFromDepth* a, void* stream) {
    EDE_REQUIRE(a && a->depth && a->vertices && a->uv && a->indices, "ede_mesh_from_depth: null argument");
    EDE_REQUIRE(a->h > 0 && a->w > 0 && a->density >= 0 && a->density <= 12 && a->depth_stride >= 1,
                "ede_mesh_from_depth: bad shape (density 0..12)");
    const int n = (1 << a->density) + 1;
    ede::mesh_from_depth_kernel<<<ede::ceil_div(n * n, 256), 256, 0, (cudaStream_t)stream>>>(*a);
    EDE_LAUNCH_CHECK("mesh_from_depth_kernel");
    return EDE_OK;
}

int64_t ede_render_workspace_bytes(const EdeRender* a) {
    if (a == nullptr || a->n_frames <= 0 || a->n_vertices <= 0 || a->n_triangles <= 0 || a->width <= 0 || a->height <= 0) return -1;
    const size_t vtx = ede::align256((size_t)a->n_frames * a->n_vertices * sizeof(ede::RVertex));
    const size_t key = ede::align256((size_t)a->n_frames * a->width * a->height * 8);
    const size_t big = ede::align256(256 + (size_t)a->n_frames * a->n_triangles * sizeof(int2));
    return (int64_t)(vtx + key + big);
}

int ede_render_mesh(const EdeRender* a, void* stream) {
    EDE_REQUIRE(a && a->vertices && a->uv && a->indices && a->mvp && a->texture && a->frames && a->workspace,
                "ede_render_mesh: null argument");
    EDE_REQUIRE(a->n_frames > 0 && a->n_frames <= 65535 && a->n_vertices > 0 && a->n_triangles > 0 && a->width > 0 &&
                    a->height > 0 && a->width <= 32768 && a->height <= 32768 && a->tex_w > 0 && a->tex_h > 0,
                "ede_render_mesh: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = reinterpret_cast<char*>(a->workspace);
    ede::RVertex* vtx = reinterpret_cast<ede::RVertex*>(ws);
    ws += ede::align256((size_t)a->n_frames * a->n_vertices * sizeof(ede::RVertex));
    unsigned long long* key = reinterpret_cast<unsigned long long*>(ws);
    const long long nkey = (long long)a->n_frames * a->width * a->height;
    ws += ede::align256((size_t)nkey * 8);
    int* big_count = reinterpret_cast<int*>(ws);
    int2* big_list = reinterpret_cast<int2*>(ws + 256);
    ede::render_vertex_kernel<<<dim3(ede::ceil_div(a->n_vertices, 256), a->n_frames), 256, 0, st>>>(*a, vtx);
    EDE_LAUNCH_CHECK("render_vertex_kernel");
    ede::render_clear_kernel<<<(int)std::min<long long>((nkey + 255) / 256, 148 * 16), 256, 0, st>>>(key, nkey, big_count);
    EDE_LAUNCH_CHECK("render_clear_kernel");
    ede::render_raster_kernel<<<dim3(ede::ceil_div(a->n_triangles, 256), a->n_frames), 256, 0, st>>>(*a, vtx, key, big_count, big_list);
    EDE_LAUNCH_CHECK("render_raster_kernel");
    ede::render_raster_big_kernel<<<148 * 4, 256, 0, st>>>(*a, vtx, key, big_count, big_list);
    EDE_LAUNCH_CHECK("render_raster_big_kernel");
    const long long npx = (long long)a->width * a->height;
    ede::render_shade_kernel<<<dim3((unsigned)((npx + 255) / 256), a->n_frames), 256, 0, st>>>(*a, vtx, key);
    EDE_LAUNCH_CHECK("render_shade_kernel");
    return EDE_OK;
}

}  // extern "C"
