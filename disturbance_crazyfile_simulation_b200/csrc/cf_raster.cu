// cf_raster.cu - headless rasteriser for capture_image() / render('rgb_array') (sm_100a).
//
// What it replaces: BulletClient.getCameraImage(1920, 1080, view, proj, renderer=ER_BULLET_HARDWARE_OPENGL) in
// envs/base.py:520-544 (capture_image, one call per control step of a validation episode, iwpg.py:513-536) and
// envs/base.py:379-416 (render).  No graphics API: frames are written into a device buffer.
//
// Four launches for all frames of a batch:
//   k_frame_ctx   one thread per frame: eye, rotation matrix, packed sky / ground colours, the screen-space bound of the
//                 drone's bounding sphere;
//   k_background  every pixel gets the sky or the ground checker from its view ray in closed form: a CTA owns 128 x 32
//                 pixels, a thread 4 columns x 4 rows, so a warp writes 384 contiguous bytes of a row and everything
//                 that depends on the column only is computed once; no shared memory, no barrier;
//   k_tri_setup   one thread per (frame, triangle): body -> world -> camera -> pixel coordinates, flat Lambert colour,
//                 a 48-byte record + a bounding box;
//   k_raster      only over the 32 x 8 pixel blocks the bounding sphere can touch (for follow cameras the grid is a
//                 small window hung at the sphere's screen bound): one pixel per thread, a warp owns a compact 8 x 4
//                 patch, scans the frame's bounding boxes 32 at a time and tests its pixel centres against the
//                 triangles that reach the patch (edge functions, 1/z interpolated with the barycentrics, nearest wins,
//                 equal depth -> lower triangle index, so the result does not depend on any order of evaluation).
// Bound: HBM writes, 3 bytes per pixel (6.2 MB per 1920x1080 frame); the triangle data (a few hundred KB) lives in L2.
// Arithmetic: float32 through the non-contracting intrinsics, mirrored operation by operation by
// oracle/raster_oracle.py (NumPy float32), which the parity tests compare with byte for byte.
#include <cuda_runtime.h>
#include <math.h>
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
    void* rec = nullptr;         // [rec_frames][n_tris] TriRec + [rec_frames][n_tris] float4 boxes, written by k_tri_setup
    int rec_frames = 0;
    // the drone pass (k_tri_setup, k_raster) runs on a side stream beside k_background: the two write disjoint pixel blocks
    cudaStream_t side = nullptr;
    cudaEvent_t ev_ctx = nullptr, ev_done = nullptr;
};

namespace {

constexpr int BW = 128, BH = 8, NT = 256;

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
    float ihw, ihh;              // 2 / width, 2 / height
    int bx0, bx1, by0, by1;      // pixel bounds of the triangles that survive set-up (k_tri_setup: atomicMin / atomicMax)
    int wx, wy;                  // first block of the drone window (bounded mode): floor((cx - rad) / 32), floor((cy - rad) / 8)
    int dx0, dx1, dy0, dy1;      // 32 x 8 pixel blocks [dx0, dx1] x [dy0, dy1] that belong to the drone pass (empty: dx0 > dx1)
};
constexpr int RW = 32, RH = 8;
struct DroneWin { int on, bounded, gx, gy; };

struct TriRec { float4 a, b, c; };   // (x0 y0 x1 y1) (x2 y2 1/area 1/z0) (1/z1 1/z2 rgb idx)

// one thread per frame: everything that depends on the pose only
__global__ void k_frame_ctx(const __grid_constant__ cf_camera cam, float radius, const float* __restrict__ poses, int n_frames,
                            const DroneWin win, FrameCtx* __restrict__ ctx) {
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
    fc.ihw = D(2.f, (float)cam.width);
    fc.ihh = D(2.f, (float)cam.height);
    fc.bx0 = fc.by0 = 0x7fffffff;
    fc.bx1 = fc.by1 = -0x7fffffff - 1;
    // Blocks of the drone pass: k_raster rewrites exactly these whole, k_background leaves exactly these alone, so the two run
    // side by side.  The blocks the screen bound of the bounding sphere touches plus one block of slack on every side, cut to
    // the window of blocks k_raster is launched over (bounded mode) and to the frame.
    const int nbx = (cam.width + RW - 1) / RW, nby = (cam.height + RH - 1) / RH;
    fc.wx = (int)floorf((fc.cx - fc.rad) / (float)RW);
    fc.wy = (int)floorf((fc.cy - fc.rad) / (float)RH);
    int x0 = 0, x1 = nbx - 1, y0 = 0, y1 = nby - 1;
    if (fc.rad >= 0.f) {
        x0 = max(x0, (int)floorf((fc.cx - fc.rad) / (float)RW) - 1); x1 = min(x1, (int)floorf((fc.cx + fc.rad) / (float)RW) + 1);
        y0 = max(y0, (int)floorf((fc.cy - fc.rad) / (float)RH) - 1); y1 = min(y1, (int)floorf((fc.cy + fc.rad) / (float)RH) + 1);
    }
    if (win.bounded) {
        x0 = max(x0, fc.wx); x1 = min(x1, fc.wx + win.gx - 1);
        y0 = max(y0, fc.wy); y1 = min(y1, fc.wy + win.gy - 1);
    }
    if (!win.on) { x0 = 1; x1 = 0; }
    fc.dx0 = x0; fc.dx1 = x1; fc.dy0 = y0; fc.dy1 = y1;
    ctx[frame] = fc;
}

// Background and ground of every pixel: a CTA owns 128 x 32 pixels of one frame; a thread owns 4 columns x 4 rows
// (rows ty, ty + 8, ty + 16, ty + 24 of the block), so everything that depends on the column only is computed once and
// a warp still writes 384 contiguous bytes of a row.  No shared memory, no barrier.
constexpr int GH = 32, GROWS = GH / BH;
__device__ __forceinline__ float keep(float v) { asm volatile("" : "+f"(v)); return v; }   // held in a register, not recomputed
template <bool FLAT>                                        // FLAT: camera without roll (right_z == 0)
__global__ void __launch_bounds__(NT) k_background(const __grid_constant__ cf_camera cam, const FrameCtx* __restrict__ ctx,
                                                   uint8_t* __restrict__ out) {
    const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
    const int frame = blockIdx.z, px0 = blockIdx.x * BW + 4 * tx, by0 = blockIdx.y * GH;
    const int W = cam.width, H = cam.height;
    if (px0 >= W) return;
    const FrameCtx* c = ctx + frame;
    const int dy0 = __ldg(&c->dy0), dy1 = __ldg(&c->dy1);
    const bool in_cols = (px0 >> 5) >= __ldg(&c->dx0) && (px0 >> 5) <= __ldg(&c->dx1);
    const float ex = keep(__ldg(&c->eye[0])), ey = keep(__ldg(&c->eye[1])), ez = __ldg(&c->eye[2]);
    const float ihw = __ldg(&c->ihw), ihh = __ldg(&c->ihh);
    const uint32_t bg = __ldg(&c->bg), ga = __ldg(&c->ga), gb = __ldg(&c->gb);
    const bool above = ez > 0.f;
    const float gh = keep(cam.ground_half), ick = keep(cam.inv_checker);
    const float pxc0 = A((float)px0, 0.5f);
    float rx[4], ry[4], rz[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float txv = M(Sb(M(A(pxc0, (float)j), ihw), 1.f), cam.tan_x);
        rx[j] = keep(M(cam.right[0], txv)); ry[j] = keep(M(cam.right[1], txv));
        rz[j] = FLAT ? 0.f : keep(M(cam.right[2], txv));
    }
    uint8_t* obase = out + ((size_t)frame * H) * (size_t)W * 3 + (size_t)px0 * 3;
#pragma unroll
    for (int r = 0; r < GROWS; ++r) {
        const int py = by0 + ty + BH * r;
        if (py >= H) break;
        if (in_cols && (py >> 3) >= dy0 && (py >> 3) <= dy1) continue;   // the drone pass (k_raster) writes this 32 x 8 block
        const float tyv = M(Sb(1.f, M(A((float)py, 0.5f), ihh)), cam.tan_y);
        const float cxr = A(cam.fwd[0], M(cam.up[0], tyv)), cyr = A(cam.fwd[1], M(cam.up[1], tyv)), czr = A(cam.fwd[2], M(cam.up[2], tyv));
        uint32_t col[4] = {bg, bg, bg, bg};
        if (FLAT) {                                         // the whole row is ground or sky: one division per row, selects per pixel
            if (above && czr < 0.f) {
                const float t = D(ez, -czr);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float hx = A(ex, M(t, A(cxr, rx[j]))), hy = A(ey, M(t, A(cyr, ry[j])));
                    const uint32_t g = ((__float2int_rd(M(hx, ick)) + __float2int_rd(M(hy, ick))) & 1) ? gb : ga;
                    col[j] = (fabsf(hx) <= gh && fabsf(hy) <= gh) ? g : bg;
                }
            }
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float dz = A(czr, rz[j]);
                if (dz < 0.f && above) {
                    const float t = D(ez, -dz);
                    const float hx = A(ex, M(t, A(cxr, rx[j]))), hy = A(ey, M(t, A(cyr, ry[j])));
                    if (fabsf(hx) <= gh && fabsf(hy) <= gh)
                        col[j] = ((__float2int_rd(M(hx, ick)) + __float2int_rd(M(hy, ick))) & 1) ? gb : ga;
                }
            }
        }
        uint32_t* o = reinterpret_cast<uint32_t*>(obase + (size_t)py * (size_t)W * 3);
        o[0] = col[0] | (col[1] << 24);
        o[1] = (col[1] >> 8) | (col[2] << 16);
        o[2] = (col[2] >> 16) | (col[3] << 8);
    }
}

// ---- drone pass ------------------------------------------------------------------------------------------------
// k_tri_setup: one thread per (frame, triangle): body -> world -> camera -> pixel coordinates, flat Lambert colour;
// writes a 48-byte record and a bounding box (an empty box for a triangle that is dropped).
__global__ void __launch_bounds__(NT) k_tri_setup(const float* __restrict__ tris, const float* __restrict__ rgb, int n_tris,
                                                  const __grid_constant__ cf_camera cam, FrameCtx* __restrict__ ctx,
                                                  TriRec* __restrict__ rec, float4* __restrict__ bbs) {
    const int frame = blockIdx.y;
    const bool valid = (int)(blockIdx.x * NT + threadIdx.x) < n_tris;       // the tail threads redo the last triangle and write nothing
    const int ti = valid ? blockIdx.x * NT + threadIdx.x : n_tris - 1;
    FrameCtx* c = ctx + frame;
    float pos[3], eye[3], R[9];
#pragma unroll
    for (int k = 0; k < 3; ++k) { pos[k] = __ldg(&c->pos[k]); eye[k] = __ldg(&c->eye[k]); }
#pragma unroll
    for (int k = 0; k < 9; ++k) R[k] = __ldg(&c->R[k]);
    const float half_w = M((float)cam.width, 0.5f), half_h = M((float)cam.height, 0.5f);      // == W / 2, H / 2
    const float* tv = tris + 9 * (size_t)ti;
    float wv[3][3], sx[3], sy[3], iz[3];
    bool ok = true;
#pragma unroll
    for (int v = 0; v < 3; ++v) {
        const float bx = tv[3 * v], by = tv[3 * v + 1], bz = tv[3 * v + 2];
#pragma unroll
        for (int k = 0; k < 3; ++k) wv[v][k] = A(pos[k], dot3(bx, by, bz, R + 3 * k));
        float pc[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) pc[k] = Sb(wv[v][k], eye[k]);
        const float xc = dot3(pc, cam.right), yc = dot3(pc, cam.up), zc = dot3(pc, cam.fwd);
        ok = ok && zc >= cam.near_plane;
        sx[v] = M(A(D(D(xc, zc), cam.tan_x), 1.f), half_w);
        sy[v] = M(Sb(1.f, D(D(yc, zc), cam.tan_y)), half_h);
        iz[v] = D(1.f, zc);
    }
    const float area = Sb(M(Sb(sx[2], sx[0]), Sb(sy[1], sy[0])), M(Sb(sy[2], sy[0]), Sb(sx[1], sx[0])));
    ok = ok && area != 0.f;
    float e1[3], e2[3], ev[3], n[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) { e1[k] = Sb(wv[1][k], wv[0][k]); e2[k] = Sb(wv[2][k], wv[0][k]); ev[k] = Sb(eye[k], wv[0][k]); }
    n[0] = Sb(M(e1[1], e2[2]), M(e1[2], e2[1]));
    n[1] = Sb(M(e1[2], e2[0]), M(e1[0], e2[2]));
    n[2] = Sb(M(e1[0], e2[1]), M(e1[1], e2[0]));
    const float len = __fsqrt_rn(dot3(n, n));
    ok = ok && len > 0.f;
    const size_t o = (size_t)frame * n_tris + ti;
    const float mnx = fminf(sx[0], fminf(sx[1], sx[2])), mxx = fmaxf(sx[0], fmaxf(sx[1], sx[2]));
    const float mny = fminf(sy[0], fminf(sy[1], sy[2])), mxy = fmaxf(sy[0], fmaxf(sy[1], sy[2]));
    {                                                       // pixel bounds of the frame's drone: warp reduction, 4 atomics per warp
        const float lim = 1.0e9f;
        const bool in = ok && valid;
        const int a0 = __reduce_min_sync(0xffffffffu, in ? (int)floorf(fmaxf(mnx, -lim)) : 0x7fffffff);
        const int a1 = __reduce_max_sync(0xffffffffu, in ? (int)ceilf(fminf(mxx, lim)) : -0x7fffffff - 1);
        const int a2 = __reduce_min_sync(0xffffffffu, in ? (int)floorf(fmaxf(mny, -lim)) : 0x7fffffff);
        const int a3 = __reduce_max_sync(0xffffffffu, in ? (int)ceilf(fminf(mxy, lim)) : -0x7fffffff - 1);
        if ((threadIdx.x & 31) == 0 && a0 != 0x7fffffff) {
            atomicMin(&c->bx0, a0); atomicMax(&c->bx1, a1); atomicMin(&c->by0, a2); atomicMax(&c->by1, a3);
        }
    }
    if (!valid) return;
    if (!ok) {                                              // empty box: no patch ever picks it
        bbs[o] = make_float4(3.0e38f, -3.0e38f, 3.0e38f, -3.0e38f);
        return;
    }
    float ndl = D(dot3(n, cam.light), len);
    if (dot3(n, ev) < 0.f) ndl = -ndl;                      // two-sided: the normal is turned towards the eye
    const float sh = A(cam.ambient, M(cam.diffuse, fmaxf(ndl, 0.f)));
    const float* cr = rgb + 3 * (size_t)ti;
    const uint32_t pk = pack_rgb(M(cr[0], sh), M(cr[1], sh), M(cr[2], sh));
    rec[o].a = make_float4(sx[0], sy[0], sx[1], sy[1]);
    rec[o].b = make_float4(sx[2], sy[2], D(1.f, area), iz[0]);
    rec[o].c = make_float4(iz[1], iz[2], __uint_as_float(pk), __int_as_float(ti));
    bbs[o] = make_float4(mnx, mxx, mny, mxy);
}

// k_raster: runs behind k_background over the pixel blocks the drone's bounding sphere can touch and rewrites them whole.
// A CTA owns 32 x 8 pixels, one pixel per thread, a warp a compact 8 x 4 patch: the drone is ~50 pixels wide with hundreds
// of triangles, so small patches spread the triangle walk over many warps.  No shared memory, no barrier: a warp scans the
// frame's bounding boxes 32 at a time (one per lane) and walks only the triangles that reach its patch.
// bounded != 0: the grid is a window of blocks hung at the sphere's screen bound (follow cameras, where the host knows its
// size); else the grid covers the frame and untouched CTAs leave at once.
__global__ void __launch_bounds__(NT) k_raster(int n_tris, const __grid_constant__ cf_camera cam, const FrameCtx* __restrict__ ctx,
                                               const TriRec* __restrict__ rec, const float4* __restrict__ bbs, const DroneWin win,
                                               uint8_t* __restrict__ out) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wx0 = 8 * (warp & 3), wy0 = 4 * (warp >> 2);
    const int frame = blockIdx.z;
    const int W = cam.width, H = cam.height;
    const FrameCtx* c = ctx + frame;
    int tbx = blockIdx.x, tby = blockIdx.y;
    if (win.bounded) { tbx += __ldg(&c->wx); tby += __ldg(&c->wy); }
    if (tbx < __ldg(&c->dx0) || tbx > __ldg(&c->dx1) || tby < __ldg(&c->dy0) || tby > __ldg(&c->dy1)) return;   // k_background's block
    // the pixel bounds k_tri_setup found (one pixel of slack): a block or patch outside them skips the triangle scan and
    // writes the background only (it still writes: k_background left the block alone)
    const int qx0 = c->bx0 - 1, qx1 = c->bx1 + 1, qy0 = c->by0 - 1, qy1 = c->by1 + 1;
    const bool live = !(tbx * RW + wx0 > qx1 || tbx * RW + wx0 + 8 < qx0 || tby * RH + wy0 > qy1 || tby * RH + wy0 + 4 < qy0);
    const int px = tbx * RW + wx0 + (lane & 7), py = tby * RH + wy0 + (lane >> 3);

    // background / ground of this pixel, as in k_background, keeping the depth of the ground hit
    const float ex = __ldg(&c->eye[0]), ey = __ldg(&c->eye[1]), ez = __ldg(&c->eye[2]);
    const float pxc = A((float)px, 0.5f), pyc = A((float)py, 0.5f);
    const float txv = M(Sb(M(pxc, __ldg(&c->ihw)), 1.f), cam.tan_x), tyv = M(Sb(1.f, M(pyc, __ldg(&c->ihh))), cam.tan_y);
    const float dx = A(A(cam.fwd[0], M(cam.up[0], tyv)), M(cam.right[0], txv));
    const float dy = A(A(cam.fwd[1], M(cam.up[1], tyv)), M(cam.right[1], txv));
    const float dz = A(A(cam.fwd[2], M(cam.up[2], tyv)), M(cam.right[2], txv));
    uint32_t col = __ldg(&c->bg);
    float best = 0.f;                                       // 1 / depth of the nearest surface so far (0: none)
    int best_id = 0x7fffffff;
    if (dz < 0.f && ez > 0.f) {
        const float t = D(ez, -dz);
        const float hx = A(ex, M(t, dx)), hy = A(ey, M(t, dy));
        if (fabsf(hx) <= cam.ground_half && fabsf(hy) <= cam.ground_half) {
            col = ((__float2int_rd(M(hx, cam.inv_checker)) + __float2int_rd(M(hy, cam.inv_checker))) & 1) ? __ldg(&c->gb) : __ldg(&c->ga);
            best = D(1.f, t);
        }
    }

    // pixel centres of the warp's patch span [wl + 1, wr - 1] x [wt + 1, wb - 1]: one pixel of slack around them
    const float wl = (float)(tbx * RW + wx0) - 0.5f, wr = (float)(tbx * RW + wx0 + 8) + 0.5f;
    const float wt = (float)(tby * RH + wy0) - 0.5f, wb = (float)(tby * RH + wy0 + 4) + 0.5f;
    const float4* fb = bbs + (size_t)frame * n_tris;
    const TriRec* fr = rec + (size_t)frame * n_tris;
    for (int r0 = 0; live && r0 < n_tris; r0 += 32) {
        bool hit = false;
        if (r0 + lane < n_tris) {
            const float4 bb = __ldg(fb + r0 + lane);
            hit = !(bb.x > wr || bb.y < wl || bb.z > wb || bb.w < wt);
        }
        unsigned todo = __ballot_sync(0xffffffffu, hit);
        if (todo == 0u) continue;
        // every lane with a hit fetches its own record (all loads in flight together); the warp then takes the records
        // one by one out of the owners' registers, so no global-memory latency sits inside the serial walk
        float4 ma = make_float4(0.f, 0.f, 0.f, 0.f), mb = ma, mc = ma;
        if (hit) { const TriRec* q = fr + r0 + lane; ma = __ldg(&q->a); mb = __ldg(&q->b); mc = __ldg(&q->c); }
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const float x0 = __shfl_sync(0xffffffffu, ma.x, src), y0 = __shfl_sync(0xffffffffu, ma.y, src);
            const float x1 = __shfl_sync(0xffffffffu, ma.z, src), y1 = __shfl_sync(0xffffffffu, ma.w, src);
            const float x2 = __shfl_sync(0xffffffffu, mb.x, src), y2 = __shfl_sync(0xffffffffu, mb.y, src);
            const float ia = __shfl_sync(0xffffffffu, mb.z, src), iz0 = __shfl_sync(0xffffffffu, mb.w, src);
            const float iz1 = __shfl_sync(0xffffffffu, mc.x, src), iz2 = __shfl_sync(0xffffffffu, mc.y, src);
            const float b0 = M(Sb(M(Sb(pxc, x1), Sb(y2, y1)), M(Sb(pyc, y1), Sb(x2, x1))), ia);
            const float b1 = M(Sb(M(Sb(pxc, x2), Sb(y0, y2)), M(Sb(pyc, y2), Sb(x0, x2))), ia);
            const float b2 = M(Sb(M(Sb(pxc, x0), Sb(y1, y0)), M(Sb(pyc, y0), Sb(x1, x0))), ia);
            const bool inside = b0 >= 0.f && b1 >= 0.f && b2 >= 0.f;
            if (__any_sync(0xffffffffu, inside)) {
                const uint32_t pk = __shfl_sync(0xffffffffu, __float_as_uint(mc.z), src);
                const int id = __shfl_sync(0xffffffffu, __float_as_int(mc.w), src);
                const float izp = A(A(M(b0, iz0), M(b1, iz1)), M(b2, iz2));
                if (inside && (izp > best || (izp == best && id < best_id))) { best = izp; best_id = id; col = pk; }
            }
        }
    }
    if (px < W && py < H) {
        uint8_t* o = out + ((size_t)frame * H + py) * (size_t)W * 3 + (size_t)px * 3;
        o[0] = (uint8_t)col; o[1] = (uint8_t)(col >> 8); o[2] = (uint8_t)(col >> 16);
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
    if (cudaStreamCreateWithFlags(&r->side, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&r->ev_ctx, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&r->ev_done, cudaEventDisableTiming) != cudaSuccess) {
        cf_raster_destroy(r);
        return fail(CF_ERR_CUDA, "cf_raster_create", "could not create the side stream");
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
    cudaFree(r->rec);
    if (r->ev_ctx) cudaEventDestroy(r->ev_ctx);
    if (r->ev_done) cudaEventDestroy(r->ev_done);
    if (r->side) cudaStreamDestroy(r->side);
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
    // frames are processed in batches so that the per-(frame, triangle) records stay within ~256 MB
    const size_t per_frame = (size_t)(r->n_tris > 0 ? r->n_tris : 1) * (sizeof(TriRec) + sizeof(float4));
    int batch = (int)((size_t)256 << 20) / (int)(per_frame < ((size_t)256 << 20) ? per_frame : ((size_t)256 << 20));
    if (batch < 1) batch = 1;
    if (batch > n_frames) batch = n_frames;
    if (batch > r->ctx_frames) {                            // grows on demand (cudaFree waits for launches still using it)
        cudaFree(r->ctx);
        cudaFree(r->rec);
        r->ctx = r->rec = nullptr; r->ctx_frames = r->rec_frames = 0;
        if (cudaMalloc(&r->ctx, (size_t)batch * sizeof(FrameCtx)) != cudaSuccess || cudaMalloc(&r->rec, (size_t)batch * per_frame) != cudaSuccess)
            return fail(CF_ERR_ALLOC, "cf_raster_render_dev", "out of device memory");
        r->ctx_frames = r->rec_frames = batch;
    }
    FrameCtx* ctx = static_cast<FrameCtx*>(r->ctx);
    TriRec* rec = static_cast<TriRec*>(r->rec);
    float4* bbs = reinterpret_cast<float4*>(rec + (size_t)r->ctx_frames * (r->n_tris > 0 ? r->n_tris : 1));
    cudaStream_t st = (cudaStream_t)stream;
    // follow camera: the drone sits at the distance |eye_offset| on the view axis, so the size of its screen bound is
    // known here (k_frame_ctx's formula with |x/z| = |y/z| = 0, plus slack for its float rounding)
    const double dist = sqrt((double)cam->eye_offset[0] * cam->eye_offset[0] + (double)cam->eye_offset[1] * cam->eye_offset[1] +
                             (double)cam->eye_offset[2] * cam->eye_offset[2]);
    dim3 grid((cam->width + RW - 1) / RW, (cam->height + RH - 1) / RH, 1);
    int bounded = 0;
    if (cam->follow && dist - r->radius > cam->near_plane + 0.01) {
        const double focal = fmax(0.5 * cam->width / cam->tan_x, 0.5 * cam->height / cam->tan_y);
        const double rad = r->radius / (dist - r->radius) * 1.01 * focal + 3.0;
        const unsigned nx = (unsigned)floor(2.0 * rad / RW) + 2, ny = (unsigned)floor(2.0 * rad / RH) + 2;
        if (nx < grid.x || ny < grid.y) {
            bounded = 1;
            grid.x = nx < grid.x ? nx : grid.x;
            grid.y = ny < grid.y ? ny : grid.y;
        }
    }
    for (int f0 = 0; f0 < n_frames; f0 += batch) {
        const int nf = n_frames - f0 < batch ? n_frames - f0 : batch;
        uint8_t* o = rgb_dev + (size_t)f0 * cam->height * cam->width * 3;
        const DroneWin win{r->n_tris > 0 ? 1 : 0, bounded, (int)grid.x, (int)grid.y};
        k_frame_ctx<<<(nf + 127) / 128, 128, 0, st>>>(*cam, r->radius, poses_dev + 7 * (size_t)f0, nf, win, ctx);
        if (r->n_tris > 0) {                                // drone pass on the side stream, beside the background pass
            RCU(cudaEventRecord(r->ev_ctx, st));
            RCU(cudaStreamWaitEvent(r->side, r->ev_ctx, 0));
            k_tri_setup<<<dim3((r->n_tris + NT - 1) / NT, nf), NT, 0, r->side>>>(r->tris, r->rgb, r->n_tris, *cam, ctx, rec, bbs);
            grid.z = nf;
            k_raster<<<grid, NT, 0, r->side>>>(r->n_tris, *cam, ctx, rec, bbs, win, o);
            RCU(cudaEventRecord(r->ev_done, r->side));
        }
        dim3 gbg((cam->width + BW - 1) / BW, (cam->height + GH - 1) / GH, nf);
        if (cam->right[2] == 0.f) k_background<true><<<gbg, NT, 0, st>>>(*cam, ctx, o);
        else k_background<false><<<gbg, NT, 0, st>>>(*cam, ctx, o);
        if (r->n_tris > 0) RCU(cudaStreamWaitEvent(st, r->ev_done, 0));
    }
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
