// forward.cu -- the `Rasterise` forward pass as sm_100a CUDA kernels.
//
// Replaces RasteriseOpGpu::Compute (csrc/rasterise_egl.cpp:284-514): the per-element GL draw
// loop (:440-458), the background upload and pixel download kernels
// (csrc/rasterise_egl.cu:16-129) and the three device-to-device mesh copies (:365-367).
// One launch per stage covers the whole batch; nothing is serial over batch elements.
//
// Stages (all on the caller's stream; DESIGN.md section 3 has the table):
//   S0  clear       keys := EMPTY, counters := 0
//   S0b project     (setup_raster.cu) one record per vertex
//   S1a order       (depth_order.cu) faces split into "first" (hinted or nearest) and "rest"
//   S1  setup       (setup_raster.cu) lane = face set-up, then lane = scan line; 64-bit atomicMin on the
//                   key buffer; the second launch tests faces and 8-line chunks against the depth pyramid
//   S2a bin_count   1 warp / large triangle: exact tile/triangle overlap test, per-tile counters
//   S2b bin_alloc   1 thread / tile: reserve a contiguous, 16-byte aligned queue segment
//   S2c bin_fill    1 warp / large triangle: write face references into the tile queues
//   S3  tile_raster 1 CTA / 16x16 tile: queue segment -> shared memory (bulk async copy),
//                   per-entry setup into shared memory, per-pixel register min of the key
//   S3s small_raster meshes of <= 32 faces: one launch instead of S0, S1, S2, S3
//   S4  resolve     1 thread / pixel: key -> face -> perspective-correct interpolation ->
//                   pixels (+ face_id, gbuffer), background where uncovered
#include "forward_common.cuh"

namespace dirt {

// Exact conservative test: can any pixel centre of tile (tx, ty) be covered by t?
__device__ __forceinline__ bool tile_overlaps(const TriSetup &t, const long long *sx, const long long *sy,
                                              int tx, int ty, int W, int H)
{
    int px0 = max(tx * kTile, t.xmin), px1 = min(tx * kTile + kTile - 1, t.xmax);
    int py0 = max(ty * kTile, t.ymin), py1 = min(ty * kTile + kTile - 1, t.ymax);
    if (px0 > px1 || py0 > py1) return false;
    long long e[3];
    edge_values(t, px0, py0, e[0], e[1], e[2]);
    const int bias[3] = {t.bias0, t.bias1, t.bias2};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        // maximum of the (affine) edge value over the tile's sample rectangle
        long long m = e[k] + (sx[k] > 0 ? sx[k] * (px1 - px0) : 0) + (sy[k] > 0 ? sy[k] * (py1 - py0) : 0);
        if (m + bias[k] < 0) return false;
    }
    return true;
}

// S2a / S2c.  FILL = false: count references per tile; FILL = true: write them.
template <bool FILL>
__global__ void __launch_bounds__(256) k_bin(FwdParams p)
{
    const int lane = threadIdx.x & 31;
    const int warps_per_grid = (gridDim.x * blockDim.x) >> 5;
    const uint32_t nbig = min(p.counters[kCtrBig], p.cap_big);
    const int tiles_x = p.tiles_x, tiles = p.tiles_x * p.tiles_y;
    for (uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < nbig; i += warps_per_grid) {
        uint4 raw = reinterpret_cast<const uint4 *>(p.big)[i];
        BigTri bt = *reinterpret_cast<BigTri *>(&raw);
        int b = (int)bt.elem;
        TriSetup t;
        if (!setup_from_ref(p, b, bt.ref, t)) continue;      // cannot happen: it was set up once already
        long long sx[3], sy[3];
        edge_steps(t, sx, sy);
        int nx = bt.tx1 - bt.tx0 + 1, ny = bt.ty1 - bt.ty0 + 1;
        for (int j = lane; j < nx * ny; j += 32) {
            int tx = bt.tx0 + j % nx, ty = bt.ty0 + j / nx;
            if (!tile_overlaps(t, sx, sy, tx, ty, p.W, p.H)) continue;
            size_t tile = (size_t)b * tiles + (size_t)ty * tiles_x + tx;
            if (!FILL) {
                atomicAdd(&p.tile_count[tile], 1u);
            } else {
                uint32_t k = atomicAdd(&p.tile_cursor[tile], 1u);
                unsigned long long slot = (unsigned long long)p.tile_offset[tile] + k;
                if (slot < p.cap_queue) {
                    p.queue[slot] = bt.ref;
                } else {
                    // queue full: this lane samples the tile itself (correct, slow)
                    int px0 = max(tx * kTile, t.xmin), px1 = min(tx * kTile + kTile - 1, t.xmax);
                    int py0 = max(ty * kTile, t.ymin), py1 = min(ty * kTile + kTile - 1, t.ymax);
                    raster_serial(t, bt.ref & ((1u << kSubShift) - 1u), p.keys + (size_t)b * p.H * p.W,
                                  p.W, p.H, px0, px1, py0, py1);
                }
            }
        }
    }
}

// S2b: reserve queue space per tile.  Segments start on a multiple of 4 entries (16 bytes) so
// that tile_raster can fetch them with one bulk async copy.
__global__ void __launch_bounds__(256) k_bin_alloc(FwdParams p)
{
    if (p.counters[kCtrBig] == 0) return;
    size_t n = (size_t)p.B * p.tiles_x * p.tiles_y;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t c = p.tile_count[i];
        if (c == 0) continue;
        uint32_t padded = (c + 3u) & ~3u;
        p.tile_offset[i] = atomicAdd(&p.counters[kCtrQueue], padded);
        p.active_tiles[atomicAdd(&p.counters[kCtrActiveTiles], 1u)] = (uint32_t)i;
    }
}

// ---- bulk async copy (TMA engine, non-tensor form) + mbarrier helpers --------------------
__device__ __forceinline__ uint32_t smem_u32(const void *ptr) { return (uint32_t)__cvta_generic_to_shared(ptr); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// S3: one CTA per tile, one thread per pixel.
__global__ void __launch_bounds__(kTileThreads) k_tile_raster(FwdParams p)
{
    __shared__ __align__(16) uint32_t s_ref[kChunk];
    __shared__ TriSetup s_tri[kChunk];
    __shared__ uint8_t s_ok[kChunk];
    __shared__ __align__(8) uint64_t s_bar;

    const int tiles = p.tiles_x * p.tiles_y;
    const uint32_t nactive = p.counters[kCtrActiveTiles];
    if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    __syncthreads();
    uint32_t parity = 0;
    // the CTAs walk the list of tiles that have a queue (k_bin_alloc): most scenes leave most
    // tiles without large triangles, and a CTA launch per empty tile costs more than the work
    for (uint32_t at = blockIdx.x; at < nactive; at += gridDim.x) {
    const size_t tile = p.active_tiles[at];
    const uint32_t count = p.tile_count[tile];
    const int b = (int)(tile / tiles), tl = (int)(tile - (size_t)b * tiles);
    const int ty = tl / p.tiles_x, tx = tl - ty * p.tiles_x;
    const int px = tx * kTile + (threadIdx.x & (kTile - 1)), py = ty * kTile + (threadIdx.x >> 4);
    const bool in_image = px < p.W && py < p.H;
    unsigned long long *keyp = p.keys + (size_t)b * p.H * p.W + (size_t)(p.H - 1 - py) * p.W + px;
    unsigned long long best = in_image ? *keyp : 0ull, orig = best;

    // entries past the queue capacity were sampled by bin_fill's fallback
    const unsigned long long off = p.tile_offset[tile];
    uint32_t avail = off >= p.cap_queue ? 0u : (uint32_t)min((unsigned long long)count, p.cap_queue - off);

    for (uint32_t base = 0; base < avail; base += kChunk) {
        uint32_t n = min((uint32_t)kChunk, avail - base);
        // (1) queue segment -> shared memory: one bulk async copy (16-byte granules)
        if (threadIdx.x == 0) {
            uint32_t bytes = ((n + 3u) & ~3u) * 4u;
            // the padded tail may reach past the queue's end only by whole granules we own
            mbar_expect_tx(&s_bar, bytes);
            bulk_g2s(s_ref, p.queue + off + base, bytes, &s_bar);
        }
        mbar_wait(&s_bar, parity);
        parity ^= 1u;
        // (2) one thread per entry re-derives the triangle's setup into shared memory
        if (threadIdx.x < n) {
            TriSetup t;
            bool ok = setup_from_ref(p, b, s_ref[threadIdx.x], t);
            s_ok[threadIdx.x] = ok;
            if (ok) s_tri[threadIdx.x] = t;
        }
        __syncthreads();
        // (3) every pixel walks the staged triangles; the running minimum stays in a register
        if (in_image) {
            for (uint32_t j = 0; j < n; ++j) {
                if (!s_ok[j]) continue;
                const TriSetup &t = s_tri[j];
                if (px < t.xmin || px > t.xmax || py < t.ymin || py > t.ymax) continue;
                long long e0, e1, e2;
                edge_values(t, px, py, e0, e1, e2);
                if (!covered(t, e0, e1, e2)) continue;
                unsigned long long key;
                if (depth_key(sample_depth(t, px, py), s_ref[j] & ((1u << kSubShift) - 1u), key)) best = min(best, key);
            }
        }
        __syncthreads();
    }
    if (in_image && best < orig) *keyp = best;
    }
}

// Small meshes (at most kSmallMeshFaces faces per element: the reference's own samples -- a quad, a
// cube -- and the full-screen quads of the scene ops): no key clear, no setup pass, no binning.
// blockIdx.y walks the batch; a CTA sets up every face of its element once (clipped faces as their
// fan triangles) into shared memory, then walks 16x16 tiles (listing per tile the triangles whose box
// touches it), thread = pixel, the running minimum in a register, and stores the key of EVERY pixel -- so the launch also stands in for the clear.
// Same set-up, coverage and depth functions as the other paths: same keys.
#ifndef KSMALLF
#define KSMALLF 32
#endif
constexpr int kSmallMeshFaces = KSMALLF;
__global__ void __launch_bounds__(kTileThreads) k_small_raster(FwdParams p)
{
    __shared__ TriSetup s_tri[3 * kSmallMeshFaces];
    __shared__ uint32_t s_face[3 * kSmallMeshFaces];
    __shared__ uint8_t s_list[3 * kSmallMeshFaces];
    __shared__ int s_n, s_m;
    const int tiles = p.tiles_x * p.tiles_y;
    for (int b = blockIdx.y; b < p.B; b += gridDim.y) {
        if (threadIdx.x == 0) s_n = 0;
        __syncthreads();
        if ((int)threadIdx.x < p.F) {
            const int f = threadIdx.x;
            const int32_t *fp = p.faces + (int64_t)b * p.faces_bs + 3 * (int64_t)f;
            const int i0 = __ldg(fp), i1 = __ldg(fp + 1), i2 = __ldg(fp + 2);
            if ((unsigned)i0 < (unsigned)p.V && (unsigned)i1 < (unsigned)p.V && (unsigned)i2 < (unsigned)p.V) {
                const int4 *pp = proj_ptr(p, b);
                TriSetup t;
                bool slow;
                if (setup_from_proj(__ldg(pp + i0), __ldg(pp + i1), __ldg(pp + i2), p.W, p.H, t, slow)) {
                    const int slot = atomicAdd(&s_n, 1);
                    s_tri[slot] = t;
                    s_face[slot] = (uint32_t)f;
                } else if (slow) {
                    const float4 *vp = vertex_ptr(p.vertices, p.vertices_bs, b);
                    const float4 va = __ldg(vp + i0), vb = __ldg(vp + i1), vc = __ldg(vp + i2);
                    if (classify_face(va, vb, vc) == 2) {
                        ClipVert poly[6];
                        const int n = clip_face(va, vb, vc, poly);
                        for (int s = 1; s + 1 < n; ++s) {
                            if (!setup_fan(poly, s, p.W, p.H, t)) continue;
                            const int slot = atomicAdd(&s_n, 1);
                            s_tri[slot] = t;
                            s_face[slot] = (uint32_t)f;
                        }
                    }
                }
            }
        }
        __syncthreads();
        const int n = s_n;
        unsigned long long *kb = p.keys + (size_t)b * p.H * p.W;
        // more than a handful of triangles: each tile first lists the ones whose box touches it
        const bool cull = n > 4;
        for (int tl = blockIdx.x; tl < tiles; tl += gridDim.x) {
            const int ty = tl / p.tiles_x, tx = tl - ty * p.tiles_x;
            const int px = tx * kTile + (threadIdx.x & (kTile - 1)), py = ty * kTile + (threadIdx.x >> 4);
            int m = n;
            if (cull) {
                if (threadIdx.x == 0) s_m = 0;
                __syncthreads();
                if ((int)threadIdx.x < n) {
                    const TriSetup &t = s_tri[threadIdx.x];
                    if (t.xmax >= tx * kTile && t.xmin < (tx + 1) * kTile && t.ymax >= ty * kTile && t.ymin < (ty + 1) * kTile)
                        s_list[atomicAdd(&s_m, 1)] = (uint8_t)threadIdx.x;
                }
                __syncthreads();
                m = s_m;
            }
            if (px < p.W && py < p.H) {
                unsigned long long best = kKeyEmpty;
                for (int k = 0; k < m; ++k) {
                    const int j = cull ? s_list[k] : k;
                    const TriSetup &t = s_tri[j];
                    if (px < t.xmin || px > t.xmax || py < t.ymin || py > t.ymax) continue;
                    long long e0, e1, e2;
                    edge_values(t, px, py, e0, e1, e2);
                    if (!covered(t, e0, e1, e2)) continue;
                    unsigned long long key;
                    if (depth_key(sample_depth(t, px, py), s_face[j], key)) best = min(best, key);
                }
                kb[(size_t)(p.H - 1 - py) * p.W + px] = best;
            }
            if (cull) __syncthreads();      // the list is about to be rewritten
        }
        __syncthreads();        // the staged triangles are about to be replaced
    }
}

// Barycentrics of pixel (px, py) on a face that went through the clipper (rare): finds the
// fan triangle whose fragment won the pixel and folds the fan weights back onto the face's own
// corners.  Normally exactly one fan triangle covers the pixel; when clipped corners land beyond
// the guard band their window positions are clamped, the fan can fold over itself, and several
// do -- then the winner is the one the depth test kept: the first (in fan order) whose 24-bit
// depth equals the depth in the key, which is what the sequential rule "nearest fragment, first
// on a tie" leaves.
static __device__ __noinline__ void resolve_clipped(const FwdParams &p, int b, int i0, int i1, int i2, int px, int py,
                                                    uint32_t depth, float &B0, float &B1, float &wc)
{
    const float4 *vp = vertex_ptr(p.vertices, p.vertices_bs, b);
    const float4 va = __ldg(vp + i0), vb = __ldg(vp + i1), vc = __ldg(vp + i2);
    ClipVert poly[6];
    const int n = clip_face(va, vb, vc, poly);
    TriSetup t;
    for (int s = 1; s + 1 < n; ++s) {
        if (!setup_fan(poly, s, p.W, p.H, t)) continue;
        if (px < t.xmin || px > t.xmax || py < t.ymin || py > t.ymax) continue;
        long long e0, e1, e2;
        edge_values(t, px, py, e0, e1, e2);
        if (!covered(t, e0, e1, e2)) continue;
        unsigned long long key;
        if (!depth_key(sample_depth(t, px, py), 0u, key) || (uint32_t)(key >> 32) != depth) continue;
        float l0, l1, l2, b0, b1, b2;
        screen_bary(t, e0, e1, e2, l0, l1, l2);
        perspective_weights(t, l0, l1, l2, b0, b1, b2, wc);
        const ClipVert &q0 = poly[0], &q1 = poly[s], &q2 = poly[s + 1];
        B0 = fadd(fadd(fmul(b0, q0.b0), fmul(b1, q1.b0)), fmul(b2, q2.b0));
        B1 = fadd(fadd(fmul(b0, q0.b1), fmul(b1, q1.b1)), fmul(b2, q2.b1));
        return;
    }
}

#ifndef KRES_PIX
#define KRES_PIX 8
#endif
constexpr int kResolvePix = KRES_PIX;

// S4: one thread per pixel (kResolvePix of them in turn); blockIdx.y walks the batch.
#ifndef KRES_CTAS
#define KRES_CTAS 4
#endif
template <int C>
__global__ void __launch_bounds__(256, KRES_CTAS) k_resolve(FwdParams p)
{
    const int CC = C > 0 ? C : p.C;
    const int npix = p.H * p.W;
    // visibility hint for the next call on this mesh: a face that wins a pixel is stamped with the next generation
    const uint8_t stamp = p.vis_flags ? (uint8_t)hint_next_generation(hint_generation(p)) : (uint8_t)0;
    // kResolvePix pixels per thread: a CTA that resolves only 256 pixels is launch-rate bound.
    // The gather chain key -> face -> corners is software-pipelined over those pixels: the key of
    // pixel j+2 and the corner indices of pixel j+1 are in flight while pixel j is shaded.
    for (int b = blockIdx.y; b < p.B; b += gridDim.y) {
    const int o0 = (int)(blockIdx.x * kResolvePix * blockDim.x + threadIdx.x);
    const unsigned long long *kb = p.keys + (size_t)b * npix;
    const int32_t *fb = p.faces + (int64_t)b * p.faces_bs;
    unsigned long long key_n1 = o0 < npix ? kb[o0] : kKeyEmpty;
    unsigned long long key_n2 = o0 + (int)blockDim.x < npix ? kb[o0 + blockDim.x] : kKeyEmpty;
    int n0 = 0, n1 = 0, n2 = 0;
    if (key_n1 != kKeyEmpty) {
        const int32_t *fp = fb + 3 * (int64_t)(uint32_t)key_n1;
        n0 = __ldg(fp); n1 = __ldg(fp + 1); n2 = __ldg(fp + 2);
    }
#pragma unroll 1
    for (int j = 0; j < kResolvePix; ++j) {
        const int o = o0 + j * (int)blockDim.x;
        if (o >= npix) break;
        const unsigned long long key = key_n1;
        const int i0 = n0, i1 = n1, i2 = n2;
        key_n1 = key_n2;
        {
            const int o2 = o + 2 * (int)blockDim.x;
            key_n2 = (j + 2 < kResolvePix && o2 < npix) ? kb[o2] : kKeyEmpty;
        }
        if (key_n1 != kKeyEmpty) {
            const int32_t *fp = fb + 3 * (int64_t)(uint32_t)key_n1;
            n0 = __ldg(fp); n1 = __ldg(fp + 1); n2 = __ldg(fp + 2);
        }
        const int row = o / p.W, px = o - row * p.W, py = p.H - 1 - row;
        const size_t i = (size_t)b * npix + o;
        float *out = p.pixels + i * CC;
        if (key == kKeyEmpty) {
            // the reference never clears the colour buffer: background shows through
            if (p.pixels) {       // (null: the stateless backward re-derives face_id / gbuffer only)
                const float *bg = p.background + (int64_t)b * p.background_bs + (size_t)o * CC;
                for (int c = 0; c < CC; ++c) out[c] = __ldg(bg + c);
            }
            if (p.face_id) {
                p.face_id[i] = -1;
                reinterpret_cast<float4 *>(p.gbuffer)[i] = make_float4(0.f, 0.f, 0.f, 1.0f);
            }
            continue;
        }
        const int f = (int)(uint32_t)(key & 0xFFFFFFFFull);
        const int4 *pp = proj_ptr(p, b);
        const int4 ra = __ldg(pp + i0), rb = __ldg(pp + i1), rc = __ldg(pp + i2);
        if (p.vis_flags) {
            // One byte store per run of pixels a face owns along a row, not per pixel: the lanes of a warp hold
            // neighbouring pixels, and a lane whose left neighbour (if it is here: the covered lanes of the warp)
            // has the same face leaves the store to it.  A missed or doubled store only costs time: the hint orders work.
            const unsigned here = __activemask();
            const uint32_t left = __shfl_up_sync(here, (uint32_t)f, 1);
            const int lane = threadIdx.x & 31;
            if (lane == 0 || !((here >> (lane - 1)) & 1u) || left != (uint32_t)f)
                p.vis_flags[(size_t)b * p.F + (uint32_t)f] = stamp;
        }
        float B0 = 0.f, B1 = 0.f, wc = 1.f;
        if (ra.x != kProjSlow && rb.x != kProjSlow && rc.x != kProjSlow) {
            // the pixel is known to be covered: only the sign and 1/area of the setup are needed
            TriSetup t;
            t.X0 = ra.x; t.Y0 = ra.y; t.rw0 = __int_as_float(ra.w);
            t.X1 = rb.x; t.Y1 = rb.y; t.rw1 = __int_as_float(rb.w);
            t.X2 = rc.x; t.Y2 = rc.y; t.rw2 = __int_as_float(rc.w);
            const int ex = max(t.X0, max(t.X1, t.X2)) - min(t.X0, min(t.X1, t.X2));
            const int ey = max(t.Y0, max(t.Y1, t.Y2)) - min(t.Y0, min(t.Y1, t.Y2));
            float l0, l1, l2, b2;
            if (ex < kSmallExtent && ey < kSmallExtent) {
                // every difference is below 2^15 (the sample lies inside the triangle), so the same
                // integers come out of 32-bit arithmetic and convert to the same floats
                const int cx = px * kSubpix + kSubpixHalf, cy = py * kSubpix + kSubpixHalf;
                const int area2 = (t.X1 - t.X0) * (t.Y2 - t.Y0) - (t.Y1 - t.Y0) * (t.X2 - t.X0);
                const int sg = area2 > 0 ? 1 : -1;
                t.rcp_area = fdiv(1.0f, __int2float_rn(area2 > 0 ? area2 : -area2));
                const int e0 = ((t.X2 - t.X1) * (cy - t.Y1) - (t.Y2 - t.Y1) * (cx - t.X1)) * sg;
                const int e1 = ((t.X0 - t.X2) * (cy - t.Y2) - (t.Y0 - t.Y2) * (cx - t.X2)) * sg;
                const int e2 = ((t.X1 - t.X0) * (cy - t.Y0) - (t.Y1 - t.Y0) * (cx - t.X0)) * sg;
                l0 = fmul(__int2float_rn(e0), t.rcp_area);
                l1 = fmul(__int2float_rn(e1), t.rcp_area);
                l2 = fmul(__int2float_rn(e2), t.rcp_area);
            } else {
                const long long area2 = orient2d(t.X0, t.Y0, t.X1, t.Y1, t.X2, t.Y2);
                t.sign = area2 > 0 ? 1 : -1;
                t.rcp_area = fdiv(1.0f, __ll2float_rn(area2 > 0 ? area2 : -area2));
                long long e0, e1, e2;
                edge_values(t, px, py, e0, e1, e2);
                screen_bary(t, e0, e1, e2, l0, l1, l2);
            }
            perspective_weights(t, l0, l1, l2, B0, B1, b2, wc);
        }
        // else: a corner went through the clipper (rare) -- k_resolve_clipped rewrites this pixel; calling
        // the clip path from here costs the whole kernel registers and a stack frame
        if (p.pixels) {
            const float *cp = p.vertex_colors + (int64_t)b * p.vertex_colors_bs;
            const float *c0 = cp + (int64_t)i0 * CC, *c1 = cp + (int64_t)i1 * CC, *c2 = cp + (int64_t)i2 * CC;
            for (int c = 0; c < CC; ++c) {
                // smooth varying (csrc/shaders.cpp:23,29), arranged so constant colours are exact
                const float k2 = __ldg(c2 + c);
                const float t0 = fsub(__ldg(c0 + c), k2), t1 = fsub(__ldg(c1 + c), k2);
                out[c] = fadd(fadd(k2, fmul(B0, t0)), fmul(B1, t1));
            }
        }
        if (p.face_id) {
            p.face_id[i] = f;
            // 4th component: the window depth as the 24-bit depth buffer holds it
            const float zq = fmul(__uint2float_rn((uint32_t)(key >> 32)), kDepthScaleInv);
            reinterpret_cast<float4 *>(p.gbuffer)[i] = make_float4(B0, B1, wc, zq);
        }
    }
    }
}

// S4b: the pixels won by faces that went through the clipper.  Does nothing unless k_project met a
// vertex that needs the clip path (counters[kCtrSlow]); then one thread per pixel re-derives the
// weights through the clipped fan and rewrites colour and G-buffer with the formulas of k_resolve.
__global__ void __launch_bounds__(256) k_resolve_clipped(FwdParams p)
{
    if (p.hint_hdr && blockIdx.x == 0 && threadIdx.x == 0) {
        // k_resolve has stamped this call's visible faces with the next generation (as a pass of its own over the key
        // buffer the stamping cost 0.06 ms, as a load-compare-store per pixel in k_resolve 0.14, as it is now 0.02): publish it
        *reinterpret_cast<uint4 *>(p.hint_hdr) = make_uint4(kHintMagic, (uint32_t)p.B, (uint32_t)p.F, hint_next_generation(hint_generation(p)));
    }
    if (p.counters[kCtrSlow] == 0u) return;
    const FwdParams pc = p;      // resolve_clipped takes a reference: keep the parameter block itself out of local memory
    const int npix = pc.H * pc.W, CC = pc.C;
    for (int b = 0; b < pc.B; ++b) {
        const unsigned long long *kb = pc.keys + (size_t)b * npix;
        const int4 *pp = proj_ptr(pc, b);
        for (int o = blockIdx.x * blockDim.x + threadIdx.x; o < npix; o += gridDim.x * blockDim.x) {
            const unsigned long long key = kb[o];
            if (key == kKeyEmpty) continue;
            const int f = (int)(uint32_t)(key & 0xFFFFFFFFull);
            const int32_t *fp = pc.faces + (int64_t)b * pc.faces_bs + 3 * (int64_t)f;
            const int i0 = __ldg(fp), i1 = __ldg(fp + 1), i2 = __ldg(fp + 2);
            if (__ldg(pp + i0).x != kProjSlow && __ldg(pp + i1).x != kProjSlow && __ldg(pp + i2).x != kProjSlow) continue;
            const int row = o / pc.W, px = o - row * pc.W, py = pc.H - 1 - row;
            float B0 = 0.f, B1 = 0.f, wc = 1.f;
            resolve_clipped(pc, b, i0, i1, i2, px, py, (uint32_t)(key >> 32), B0, B1, wc);
            const size_t i = (size_t)b * npix + o;
            if (pc.pixels) {
                float *out = pc.pixels + i * CC;
                const float *cp = pc.vertex_colors + (int64_t)b * pc.vertex_colors_bs;
                const float *c0 = cp + (int64_t)i0 * CC, *c1 = cp + (int64_t)i1 * CC, *c2 = cp + (int64_t)i2 * CC;
                for (int c = 0; c < CC; ++c) {
                    const float k2 = __ldg(c2 + c);
                    const float t0 = fsub(__ldg(c0 + c), k2), t1 = fsub(__ldg(c1 + c), k2);
                    out[c] = fadd(fadd(k2, fmul(B0, t0)), fmul(B1, t1));
                }
            }
            if (pc.face_id) {
                const float zq = fmul(__uint2float_rn((uint32_t)(key >> 32)), kDepthScaleInv);
                reinterpret_cast<float4 *>(pc.gbuffer)[i] = make_float4(B0, B1, wc, zq);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// Host-side launch sequence
// ---------------------------------------------------------------------------------------
static inline int grid_for(long long work, int block, int sm_count, int waves)
{
    long long g = (work + block - 1) / block;
    long long cap = (long long)sm_count * waves;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

static cudaError_t launch_resolve(const FwdParams &p, cudaStream_t stream, uint64_t *launches, StageTimer &tm)
{
    const dim3 g((unsigned)(((long long)p.H * p.W + 256 * kResolvePix - 1) / (256 * kResolvePix)), (unsigned)(p.B < 65535 ? p.B : 65535));
    tm.begin(kStageResolve, stream);
    switch (p.C) {
        case 1: k_resolve<1><<<g, 256, 0, stream>>>(p); break;
        case 3: k_resolve<3><<<g, 256, 0, stream>>>(p); break;
        case 9: k_resolve<9><<<g, 256, 0, stream>>>(p); break;
        default: k_resolve<0><<<g, 256, 0, stream>>>(p); break;
    }
    {
        // a small grid: in the usual case every CTA returns at once, and a launch of many thousands of
        // empty CTAs would cost more than the few microseconds of the launch itself
        long long gx = ((long long)p.H * p.W + 255) / 256;
        if (gx > 592) gx = 592;
        const dim3 g((unsigned)gx, 1);
        k_resolve_clipped<<<g, 256, 0, stream>>>(p);
    }
    tm.end(kStageResolve, stream);
    *launches += 2;
    return cudaGetLastError();
}

cudaError_t launch_forward(const FwdParams &p, int sm_count, cudaStream_t stream, uint64_t *launches, StageTimer &tm)
{
    cudaError_t e;
    const size_t npix_total = (size_t)p.B * p.H * p.W;
    const size_t ntiles = (size_t)p.B * p.tiles_x * p.tiles_y;
    if (p.F > 0 && p.F <= kSmallMeshFaces) {
        // small meshes: project, one raster launch that writes every key, resolve
        if ((e = cudaMemsetAsync(p.counters, 0, kCtrCount * sizeof(uint32_t), stream)) != cudaSuccess) return e;
        launch_project(p, sm_count, stream, tm);
        const int tiles = p.tiles_x * p.tiles_y;
        const int gy = p.B < 65535 ? p.B : 65535;
        int gx = (sm_count * 8 + gy - 1) / gy;
        if (gx > tiles) gx = tiles;
        if (gx < 1) gx = 1;
        tm.begin(kStageTileRaster, stream);
        k_small_raster<<<dim3((unsigned)gx, (unsigned)gy), kTileThreads, 0, stream>>>(p);
        tm.end(kStageTileRaster, stream);
        *launches += 2;
        return launch_resolve(p, stream, launches, tm);
    }
    // S0: keys := EMPTY; counters, tile_count, tile_cursor := 0 (contiguous in the workspace)
    tm.begin(kStageClear, stream);
    if ((e = cudaMemsetAsync(p.keys, 0xFF, npix_total * sizeof(unsigned long long), stream)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(p.counters, 0, p.zero_bytes, stream)) != cudaSuccess) return e;
    tm.end(kStageClear, stream);
    const long long nfaces = (long long)p.B * p.F;
    if (nfaces > 0) {
        launch_setup(p, sm_count, stream, launches, tm);
        tm.begin(kStageBinCount, stream);
        k_bin<false><<<sm_count * 4, 256, 0, stream>>>(p);
        tm.end(kStageBinCount, stream);
        tm.begin(kStageBinAlloc, stream);
        k_bin_alloc<<<grid_for((long long)ntiles, 256, sm_count, 8), 256, 0, stream>>>(p);
        tm.end(kStageBinAlloc, stream);
        tm.begin(kStageBinFill, stream);
        k_bin<true><<<sm_count * 4, 256, 0, stream>>>(p);
        tm.end(kStageBinFill, stream);
        tm.begin(kStageTileRaster, stream);
        k_tile_raster<<<grid_for((long long)ntiles * kTileThreads, kTileThreads, sm_count, 16), kTileThreads, 0, stream>>>(p);
        tm.end(kStageTileRaster, stream);
        *launches += 5;        // k_project is counted here, the setup launches in launch_setup
    }
    return launch_resolve(p, stream, launches, tm);
}

}  // namespace dirt
