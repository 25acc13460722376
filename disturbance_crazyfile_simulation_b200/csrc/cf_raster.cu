// cf_raster.cu - headless rasteriser for capture_image() / render('rgb_array') (sm_100a).
//
// What it replaces: BulletClient.getCameraImage(1920, 1080, view, proj, renderer=ER_BULLET_HARDWARE_OPENGL) in
// envs/base.py:520-544 (capture_image, one call per control step of a validation episode, iwpg.py:513-536) and
// envs/base.py:379-416 (render).  No graphics API: frames are written into a device buffer.
//
// Two launches for all frames of a batch: k_frame_ctx (one thread per frame: eye, rotation matrix, colours, the
// drone's screen-space bound) and k_raster, where a CTA owns a 128 x 8 pixel block of one frame (4 pixels per
// thread, so a warp writes 384 contiguous bytes of a row):
//   1. the frame's context is copied into shared memory;
//   2. every pixel gets the background or the ground checker from its view ray (closed form, no geometry);
//   3. only CTAs the drone's bounding sphere can touch walk the triangle soup in chunks of 256: each thread sets
//      one triangle up (body -> world -> camera -> pixel coordinates, flat Lambert colour), triangles whose bounding
//      box misses the CTA's block are dropped, the rest are compacted into shared memory and every thread tests its
//      four pixel centres against them (edge functions, 1/z interpolated with the barycentrics, nearest wins,
//      equal depth -> lower triangle index, so the result does not depend on the compaction order);
//   4. 12 bytes per thread leave as three 32-bit stores.
// Bound: HBM writes, 3 bytes per pixel (6.2 MB per 1920x1080 frame); the triangle data (a few hundred KB) lives in L2.
// Arithmetic: float32 through the non-contracting intrinsics, mirrored operation by operation by
// oracle/raster_oracle.py (NumPy float32), which the parity tests compare with byte for byte.
#include <cuda_runtime.h>
#include <stdint.h>

#include <new>

#include "../../include/cf_abi.h"
#include "cf_kernels.h"

struct cf_raster {
    int device = 0, n_tris = 0;
    float radius = 0.f;
    float* tris = nullptr;       // [n_tris][9]
    float* rgb = nullptr;        // [n_tris][3]
    void* ctx = nullptr;         // [ctx_frames] FrameCtx, written by k_frame_ctx in front of every k_raster launch
    int ctx_frames = 0;
};

namespace {

constexpr int BW = 128, BH = 8, NT = 256, CHUNK = 256;

__device__ __forceinline__ float M(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float A(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float Sb(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float D(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float dot3(const float* a, const float* b) { return A(A(M(a[0], b[0]), M(a[1], b[1])), M(a[2], b[2])); }
__device__ __forceinline__ float dot3(float ax, float ay, float az, const float* b) { return A(A(M(ax, b[0]), M(ay, b[1])), M(az, b[2])); }
__device__ __forceinline__ uint32_t to_u8(float v) {
    return (uint32_t)fminf(255.f, fmaxf(0.f, floorf(A(M(v, 255.f), 0.5f))));
}
__device__ __forceinline__ uint32_t pack_rgb(float r, float g, float b) { return to_u8(r) | (to_u8(g) << 8) | (to_u8(b) << 16); }

struct FrameCtx {
    float pos[3], R[9], eye[3];
    float cx, cy, rad;           // screen-space bound of the drone (pixels); rad < 0: whole screen
    uint32_t bg, ga, gb;         // packed colours: background, the two ground squares (already shaded)
};

struct TriRec { float4 a, b, c; };   // (x0 y0 x1 y1) (x2 y2 1/area 1/z0) (1/z1 1/z2 rgb idx)

// one thread per frame: everything that depends on the pose only
__global__ void k_frame_ctx(const __grid_constant__ cf_camera cam, float radius, const float* __restrict__ poses, int n_frames,
                            FrameCtx* __restrict__ ctx) {
    const int frame = blockIdx.x * blockDim.x + threadIdx.x;
    if (frame >= n_frames) return;
    FrameCtx fc;
    const float half_w = M((float)cam.width, 0.5f), half_h = M((float)cam.height, 0.5f);
    const float* ps = poses + 7 * (size_t)frame;
    const float x = ps[3], y = ps[4], z = ps[5], w = ps[6];
    const float xx = M(x, x), yy = M(y, y), zz = M(z, z), xy = M(x, y), xz = M(x, z), yz = M(y, z), wx = M(w, x), wy = M(w, y), wz = M(w, z);
    fc.R[0] = Sb(1.f, M(2.f, A(yy, zz))); fc.R[1] = M(2.f, Sb(xy, wz)); fc.R[2] = M(2.f, A(xz, wy));
    fc.R[3] = M(2.f, A(xy, wz)); fc.R[4] = Sb(1.f, M(2.f, A(xx, zz))); fc.R[5] = M(2.f, Sb(yz, wx));
    fc.R[6] = M(2.f, Sb(xz, wy)); fc.R[7] = M(2.f, A(yz, wx)); fc.R[8] = Sb(1.f, M(2.f, A(xx, yy)));
    float pc[3];
    for (int k = 0; k < 3; ++k) {
        fc.pos[k] = ps[k];
        fc.eye[k] = A(cam.follow ? ps[k] : cam.target[k], cam.eye_offset[k]);
        pc[k] = Sb(ps[k], fc.eye[k]);
    }
    const float zc = dot3(pc, cam.fwd), xc = dot3(pc, cam.right), yc = dot3(pc, cam.up);
    if (!(zc - radius > cam.near_plane)) {
        fc.rad = -1.f; fc.cx = fc.cy = 0.f;
    } else {                                                // |x'/z' - x/z| <= R (1 + |x/z|) / (z - R) for |p' - p| <= R
        const float ux = xc / zc, uy = yc / zc;
        fc.cx = (ux / cam.tan_x + 1.f) * half_w;
        fc.cy = (1.f - uy / cam.tan_y) * half_h;
        fc.rad = radius / (zc - radius) * (1.f + fabsf(ux) + fabsf(uy)) * fmaxf(half_w / cam.tan_x, half_h / cam.tan_y) + 2.f;
    }
    const float sg = A(cam.ambient, M(cam.diffuse, fmaxf(cam.light[2], 0.f)));
    fc.bg = pack_rgb(cam.background[0], cam.background[1], cam.background[2]);
    fc.ga = pack_rgb(M(cam.ground_a[0], sg), M(cam.ground_a[1], sg), M(cam.ground_a[2], sg));
    fc.gb = pack_rgb(M(cam.ground_b[0], sg), M(cam.ground_b[1], sg), M(cam.ground_b[2], sg));
    ctx[frame] = fc;
}

__global__ void __launch_bounds__(NT) k_raster(const float* __restrict__ tris, const float* __restrict__ rgb, int n_tris,
                                               const __grid_constant__ cf_camera cam, const FrameCtx* __restrict__ ctx,
                                               uint8_t* __restrict__ out) {
    __shared__ FrameCtx fc;
    __shared__ TriRec s_tri[CHUNK];
    __shared__ int s_cnt;
    const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
    const int frame = blockIdx.z, bx0 = blockIdx.x * BW, by0 = blockIdx.y * BH;
    const int W = cam.width, H = cam.height;
    const float half_w = M((float)W, 0.5f), half_h = M((float)H, 0.5f);          // == W / 2, H / 2
    static_assert(sizeof(FrameCtx) % 4 == 0 && sizeof(FrameCtx) / 4 <= NT, "context copy");
    if (tid < (int)(sizeof(FrameCtx) / 4)) reinterpret_cast<uint32_t*>(&fc)[tid] = reinterpret_cast<const uint32_t*>(ctx + frame)[tid];
    __syncthreads();

    // ---- background / ground from the view ray --------------------------------------------------
    const int px0 = bx0 + 4 * tx, py = by0 + ty;
    const float pyc = A((float)py, 0.5f);
    const float ihw = D(2.f, (float)W), ihh = D(2.f, (float)H);
    const float tyv = M(Sb(1.f, M(pyc, ihh)), cam.tan_y);
    // view ray d = (fwd + up * tyv) + right * txv: the bracket is constant along the thread's row; without camera roll
    // right_z == 0, so d_z and the ground depth t = eye_z / -d_z are too (A(c, M(0, txv)) == c: an exact rewrite)
    const float cxr = A(cam.fwd[0], M(cam.up[0], tyv)), cyr = A(cam.fwd[1], M(cam.up[1], tyv)), czr = A(cam.fwd[2], M(cam.up[2], tyv));
    const bool flat = cam.right[2] == 0.f, above = fc.eye[2] > 0.f;
    const float t_row = (flat && above && czr < 0.f) ? D(fc.eye[2], -czr) : 0.f;
    const float pxc0 = A((float)px0, 0.5f);
    uint32_t col[4];
    float tg[4];                                            // depth of the ground hit along the view axis (0: none)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float pxc = A(pxc0, (float)j);                // == (px0 + j) + 0.5 exactly (integers below 2^24)
        const float txv = M(Sb(M(pxc, ihw), 1.f), cam.tan_x);
        const float dx = A(cxr, M(cam.right[0], txv)), dy = A(cyr, M(cam.right[1], txv));
        const float dz = flat ? czr : A(czr, M(cam.right[2], txv));
        col[j] = fc.bg;
        tg[j] = 0.f;
        if (dz < 0.f && above) {
            const float t = flat ? t_row : D(fc.eye[2], -dz);
            const float hx = A(fc.eye[0], M(t, dx)), hy = A(fc.eye[1], M(t, dy));
            if (fabsf(hx) <= cam.ground_half && fabsf(hy) <= cam.ground_half) {
                const int ix = __float2int_rd(M(hx, cam.inv_checker)), iy = __float2int_rd(M(hy, cam.inv_checker));   // floor
                col[j] = ((ix + iy) & 1) ? fc.gb : fc.ga;
                tg[j] = t;
            }
        }
    }

    // ---- the drone ----------------------------------------------------------------------------------
    const bool touched = fc.rad < 0.f || ((float)bx0 <= fc.cx + fc.rad && (float)(bx0 + BW) >= fc.cx - fc.rad &&
                                          (float)by0 <= fc.cy + fc.rad && (float)(by0 + BH) >= fc.cy - fc.rad);
    if (touched) {                                          // uniform over the CTA
        float best[4];
        int best_id[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) { best[j] = tg[j] > 0.f ? D(1.f, tg[j]) : 0.f; best_id[j] = 0x7fffffff; }
        for (int c0 = 0; c0 < n_tris; c0 += CHUNK) {
            if (tid == 0) s_cnt = 0;
            __syncthreads();
            const int ti = c0 + tid;
            if (ti < n_tris) {
                const float* tv = tris + 9 * (size_t)ti;
                float wv[3][3], sx[3], sy[3], iz[3];
                bool ok = true;
#pragma unroll
                for (int v = 0; v < 3; ++v) {
                    const float bx = tv[3 * v], by = tv[3 * v + 1], bz = tv[3 * v + 2];
#pragma unroll
                    for (int k = 0; k < 3; ++k) wv[v][k] = A(fc.pos[k], dot3(bx, by, bz, fc.R + 3 * k) );
                    float pc[3];
#pragma unroll
                    for (int k = 0; k < 3; ++k) pc[k] = Sb(wv[v][k], fc.eye[k]);
                    const float xc = dot3(pc, cam.right), yc = dot3(pc, cam.up), zc = dot3(pc, cam.fwd);
                    ok = ok && zc >= cam.near_plane;
                    sx[v] = M(A(D(D(xc, zc), cam.tan_x), 1.f), half_w);
                    sy[v] = M(Sb(1.f, D(D(yc, zc), cam.tan_y)), half_h);
                    iz[v] = D(1.f, zc);
                }
                const float area = Sb(M(Sb(sx[2], sx[0]), Sb(sy[1], sy[0])), M(Sb(sy[2], sy[0]), Sb(sx[1], sx[0])));
                ok = ok && area != 0.f;
                const float mnx = fminf(sx[0], fminf(sx[1], sx[2])), mxx = fmaxf(sx[0], fmaxf(sx[1], sx[2]));
                const float mny = fminf(sy[0], fminf(sy[1], sy[2])), mxy = fmaxf(sy[0], fmaxf(sy[1], sy[2]));
                // pixel centres of the block span [bx0 + .5, bx0 + BW - .5]; one pixel of slack
                ok = ok && mnx <= (float)(bx0 + BW) + 0.5f && mxx >= (float)bx0 - 0.5f && mny <= (float)(by0 + BH) + 0.5f &&
                     mxy >= (float)by0 - 0.5f;
                if (ok) {
                    float e1[3], e2[3], ev[3], n[3];
#pragma unroll
                    for (int k = 0; k < 3; ++k) { e1[k] = Sb(wv[1][k], wv[0][k]); e2[k] = Sb(wv[2][k], wv[0][k]); ev[k] = Sb(fc.eye[k], wv[0][k]); }
                    n[0] = Sb(M(e1[1], e2[2]), M(e1[2], e2[1]));
                    n[1] = Sb(M(e1[2], e2[0]), M(e1[0], e2[2]));
                    n[2] = Sb(M(e1[0], e2[1]), M(e1[1], e2[0]));
                    const float len = __fsqrt_rn(dot3(n, n));
                    if (len > 0.f) {
                        float ndl = D(dot3(n, cam.light), len);
                        if (dot3(n, ev) < 0.f) ndl = -ndl;      // two-sided: the normal is turned towards the eye
                        const float sh = A(cam.ambient, M(cam.diffuse, fmaxf(ndl, 0.f)));
                        const float* cr = rgb + 3 * (size_t)ti;
                        const uint32_t pk = pack_rgb(M(cr[0], sh), M(cr[1], sh), M(cr[2], sh));
                        const int slot = atomicAdd(&s_cnt, 1);
                        s_tri[slot].a = make_float4(sx[0], sy[0], sx[1], sy[1]);
                        s_tri[slot].b = make_float4(sx[2], sy[2], D(1.f, area), iz[0]);
                        s_tri[slot].c = make_float4(iz[1], iz[2], __uint_as_float(pk), __int_as_float(ti));
                    }
                }
            }
            __syncthreads();
            const int cnt = s_cnt;
            for (int r = 0; r < cnt; ++r) {
                const float4 a = s_tri[r].a, b = s_tri[r].b, c = s_tri[r].c;
                const float x0 = a.x, y0 = a.y, x1 = a.z, y1 = a.w, x2 = b.x, y2 = b.y, ia = b.z;
                const float e0x = Sb(x2, x1), e0y = Sb(y2, y1), e1x = Sb(x0, x2), e1y = Sb(y0, y2), e2x = Sb(x1, x0), e2y = Sb(y1, y0);
                const float r0 = M(Sb(pyc, y1), e0x), r1 = M(Sb(pyc, y2), e1x), r2 = M(Sb(pyc, y0), e2x);
                const int id = __float_as_int(c.w);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float pxc = A(pxc0, (float)j);
                    const float b0 = M(Sb(M(Sb(pxc, x1), e0y), r0), ia);
                    const float b1 = M(Sb(M(Sb(pxc, x2), e1y), r1), ia);
                    const float b2 = M(Sb(M(Sb(pxc, x0), e2y), r2), ia);
                    if (b0 >= 0.f && b1 >= 0.f && b2 >= 0.f) {
                        const float izp = A(A(M(b0, b.w), M(b1, c.x)), M(b2, c.y));
                        if (izp > best[j] || (izp == best[j] && id < best_id[j])) {
                            best[j] = izp; best_id[j] = id; col[j] = __float_as_uint(c.z);
                        }
                    }
                }
            }
            __syncthreads();                                // the chunk is overwritten by the next one
        }
    }

    // ---- 4 pixels = 12 bytes = three words -------------------------------------------------------
    if (px0 < W && py < H) {
        uint32_t* o = reinterpret_cast<uint32_t*>(out + ((size_t)frame * H + py) * (size_t)W * 3 + (size_t)px0 * 3);
        o[0] = col[0] | (col[1] << 24);
        o[1] = (col[1] >> 8) | (col[2] << 16);
        o[2] = (col[2] >> 16) | (col[3] << 8);
    }
}

int fail(int code, const char* where, const char* what) {
    cf::set_error(where, what);
    return code;
}
#define RCU(call)                                                                       \
    do {                                                                                \
        cudaError_t e_ = (call);                                                        \
        if (e_ != cudaSuccess) return fail(CF_ERR_CUDA, #call, cudaGetErrorString(e_)); \
    } while (0)

}  // namespace

extern "C" {

int cf_raster_create(cf_raster** out, int32_t device, const float* tris_host, const float* rgb_host, int32_t n_tris,
                     float bound_radius) {
    if (!out || !tris_host || !rgb_host || n_tris < 0 || !(bound_radius >= 0.f))
        return fail(CF_ERR_INVALID, "cf_raster_create", "bad argument");
    RCU(cudaSetDevice(device));
    cf_raster* r = new (std::nothrow) cf_raster;
    if (!r) return fail(CF_ERR_ALLOC, "cf_raster_create", "out of host memory");
    r->device = device; r->n_tris = n_tris; r->radius = bound_radius;
    const size_t nt = n_tris > 0 ? (size_t)n_tris : 1;
    if (cudaMalloc(&r->tris, nt * 9 * sizeof(float)) != cudaSuccess || cudaMalloc(&r->rgb, nt * 3 * sizeof(float)) != cudaSuccess) {
        cf_raster_destroy(r);
        return fail(CF_ERR_ALLOC, "cf_raster_create", "out of device memory");
    }
    if (n_tris > 0) {
        RCU(cudaMemcpy(r->tris, tris_host, (size_t)n_tris * 9 * sizeof(float), cudaMemcpyHostToDevice));
        RCU(cudaMemcpy(r->rgb, rgb_host, (size_t)n_tris * 3 * sizeof(float), cudaMemcpyHostToDevice));
    }
    *out = r;
    return CF_OK;
}

void cf_raster_destroy(cf_raster* r) {
    if (!r) return;
    cudaSetDevice(r->device);
    cudaFree(r->tris);
    cudaFree(r->rgb);
    cudaFree(r->ctx);
    delete r;
}

int cf_raster_render_dev(cf_raster* r, const cf_camera* cam, const float* poses_dev, int32_t n_frames, uint8_t* rgb_dev,
                         void* stream) {
    if (!r || !cam || !poses_dev || !rgb_dev || n_frames < 0) return fail(CF_ERR_INVALID, "cf_raster_render_dev", "bad argument");
    if (cam->width <= 0 || cam->height <= 0 || cam->width % 4 != 0)
        return fail(CF_ERR_INVALID, "cf_raster_render_dev", "width must be a positive multiple of 4, height positive");
    if (n_frames > 65535) return fail(CF_ERR_INVALID, "cf_raster_render_dev", "at most 65535 frames per call");
    if (n_frames == 0) return CF_OK;
    RCU(cudaSetDevice(r->device));
    if (n_frames > r->ctx_frames) {                         // grows on demand (cudaFree waits for launches still using it)
        cudaFree(r->ctx);
        r->ctx = nullptr; r->ctx_frames = 0;
        if (cudaMalloc(&r->ctx, (size_t)n_frames * sizeof(FrameCtx)) != cudaSuccess)
            return fail(CF_ERR_ALLOC, "cf_raster_render_dev", "out of device memory");
        r->ctx_frames = n_frames;
    }
    FrameCtx* ctx = static_cast<FrameCtx*>(r->ctx);
    k_frame_ctx<<<(n_frames + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*cam, r->radius, poses_dev, n_frames, ctx);
    dim3 grid((cam->width + BW - 1) / BW, (cam->height + BH - 1) / BH, n_frames);
    k_raster<<<grid, NT, 0, (cudaStream_t)stream>>>(r->tris, r->rgb, r->n_tris, *cam, ctx, rgb_dev);
    RCU(cudaGetLastError());
    return CF_OK;
}

int cf_raster_render_host(cf_raster* r, const cf_camera* cam, const float* poses_host, int32_t n_frames, uint8_t* rgb_host) {
    if (!r || !cam || !poses_host || !rgb_host || n_frames < 0) return fail(CF_ERR_INVALID, "cf_raster_render_host", "bad argument");
    if (n_frames == 0) return CF_OK;
    RCU(cudaSetDevice(r->device));
    const size_t bytes = (size_t)n_frames * cam->height * cam->width * 3;
    float* dp = nullptr;
    uint8_t* di = nullptr;
    if (cudaMalloc(&dp, (size_t)n_frames * 7 * sizeof(float)) != cudaSuccess || cudaMalloc(&di, bytes) != cudaSuccess) {
        cudaFree(dp);
        return fail(CF_ERR_ALLOC, "cf_raster_render_host", "out of device memory");
    }
    int rc = CF_OK;
    cudaError_t e = cudaMemcpy(dp, poses_host, (size_t)n_frames * 7 * sizeof(float), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        rc = cf_raster_render_dev(r, cam, dp, n_frames, di, nullptr);
        if (rc == CF_OK) e = cudaMemcpy(rgb_host, di, bytes, cudaMemcpyDeviceToHost);
    }
    cudaFree(dp);
    cudaFree(di);
    if (rc != CF_OK) return rc;
    if (e != cudaSuccess) return fail(CF_ERR_CUDA, "cf_raster_render_host", cudaGetErrorString(e));
    return CF_OK;
}

}  // extern "C"
