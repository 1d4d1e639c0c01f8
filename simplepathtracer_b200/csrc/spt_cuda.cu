// spt_cuda.cu -- C ABI (include/spt_cuda.h) over the sm_100a kernels.  No CPU fallback exists:
// every entry point returns an error when CUDA is unusable.
#include "../../include/spt_cuda.h"
#include "spt_kernels.cuh"
#include "spt_wavefront.cuh"
#include "spt_streamed.cuh"
#include "spt_split.cuh"
#include "spt_grid.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace
{
thread_local std::string g_lastError;

int fail(int code, const char *what, cudaError_t ce = cudaSuccess)
{
    char buf[512];
    if (ce != cudaSuccess)
        snprintf(buf, sizeof(buf), "%s: %s (%s)", what, cudaGetErrorString(ce), cudaGetErrorName(ce));
    else
        snprintf(buf, sizeof(buf), "%s", what);
    g_lastError = buf;
    return code;
}

#define SPT_CUDA_CHECK(call)                                      \
    do                                                            \
    {                                                             \
        cudaError_t const ce_ = (call);                           \
        if (ce_ != cudaSuccess)                                   \
            return fail(SPT_ERR_CUDA, #call, ce_);                \
    } while (0)
} // namespace

// Scratch of spt_cuda_closest_hit_static (include/spt_cuda.h): rays in, answers out, for callers that must not allocate.
// In HOST code this name is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress (create_one,
// once per device, with that device current) and is what the kernel is given.  [dir3 | org3 | idx | evals], 32 bytes per ray.
static __device__ __align__(256) unsigned char g_staticScratch[size_t(SPT_STATIC_SCRATCH_RAYS) * 32u];

struct SptContext
{
    int device = 0;
    cudaDeviceProp prop{};
    cudaStream_t ownStream = nullptr;
    cudaStream_t stream = nullptr;
    cudaStream_t auxStream = nullptr;  // split rendering: the camera kernels run here, under the bounce kernel of the previous work item
    cudaStream_t auxStream2 = nullptr; // ... and the bounce kernels of every other strip here, next to the main stream's
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t evFork = nullptr, evJoin = nullptr, evPrimary[2] = {nullptr, nullptr}, evSecondary[2] = {nullptr, nullptr};
    cudaEvent_t evDone = nullptr; // multi-device: this device's accumulator is complete
    std::vector<cudaEvent_t> evKernel; // split rendering: 4 timing events per chunk (camera begin/end, bounce begin/end)
    uint32_t timedChunks = 0;
    std::mutex lock; // the reference calls its tracers from several host threads (Renderer.hpp:282-290)

    uint32_t n = 0;
    // the device scene: pointers into dSceneArena[sceneArena] (upload_scene)
    unsigned char *dSceneArena[2] = {nullptr, nullptr};
    size_t sceneArenaBytes[2] = {0, 0};
    int sceneArena = 0;
    std::vector<unsigned char> hostScenePacked;
    float4 *dGeom = nullptr;
    float4 *dShade = nullptr;
    uint8_t *dMat = nullptr;
    spt::FilterPair *dPairs = nullptr; // pre-filter records
    uint32_t *dBigIdx = nullptr;
    spt::FilterScene filter{};

    unsigned long long *dCounters = nullptr;
    SptTiming timing{};

    // scratch for the host-buffer entry points
    float4 *dAccum = nullptr;
    uint8_t *dRgb = nullptr;
    size_t accumPixels = 0;
    // split rendering: camera-form filter blob (rebuilt when the eye changes) and the first-hit staging buffers
    spt::FilterPair *dPairsCam = nullptr; // = dCamArena[camArena]: two arenas that take turns, as the scene's
    unsigned char *dCamArena[2] = {nullptr, nullptr};
    size_t camArenaBytes[2] = {0, 0};
    int camArena = 0;
    std::vector<unsigned char> hostBlob; // the general blob, kept to derive the camera form
    std::vector<double> hostGeom;        // cx, cy, cz, r per SLOT in binary64
    float camKey[16] = {NAN}; // view rows 0..2 + eye the camera records were built for, and the image size
    uint32_t camW = 0, camH = 0;
    float camEpsT = 0.f;
    float4 *dFirstHits = nullptr;
    size_t firstHitsBytes = 0;
    // SPT_FLAG_GRID (spt_grid.cuh): the uniform grid over the filterable spheres, built on first use after a scene upload
    spt::GridScene grid{};
    unsigned char *dGrid = nullptr;
    size_t gridBytes = 0;
    bool gridValid = false, gridUsable = false;
    std::vector<uint32_t> hostGrid;
    uint64_t stagingLimit = SPT_DEFAULT_STAGING_BYTES;
    // context-owned accumulator of the progressive entry points
    float4 *dProgAccum = nullptr;
    uint8_t *dProgRgb = nullptr;
    uint32_t progW = 0, progH = 0;
    void *dProbe = nullptr;
    size_t probeBytes = 0;
    void *dStaticScratch = nullptr; // device address of g_staticScratch on this context's device (cudaGetSymbolAddress)
    cudaError_t staticScratchError = cudaSuccess;
    size_t maxSmemOptin = 0;

    // multi-device contexts (spt_cuda_create_multi): `all` = this context followed by one internal context per further device
    std::vector<SptContext *> all;
    int partition = SPT_PARTITION_AUTO;
};

namespace
{
int ensure_scratch(SptContext *c, size_t pixels)
{
    if (pixels <= c->accumPixels)
        return SPT_OK;
    if (c->dAccum)
        cudaFree(c->dAccum);
    if (c->dRgb)
        cudaFree(c->dRgb);
    c->dAccum = nullptr;
    c->dRgb = nullptr;
    c->accumPixels = 0;
    SPT_CUDA_CHECK(cudaMalloc(&c->dAccum, pixels * sizeof(float4)));
    SPT_CUDA_CHECK(cudaMalloc(&c->dRgb, pixels * 3));
    c->accumPixels = pixels;
    return SPT_OK;
}

int ensure_probe(SptContext *c, size_t bytes)
{
    if (bytes <= c->probeBytes)
        return SPT_OK;
    if (c->dProbe)
        cudaFree(c->dProbe);
    c->dProbe = nullptr;
    c->probeBytes = 0;
    SPT_CUDA_CHECK(cudaMalloc(&c->dProbe, bytes));
    c->probeBytes = bytes;
    return SPT_OK;
}

float round_up_to_float(double d)
{
    float f = float(d);
    if (double(f) < d)
        f = std::nextafterf(f, INFINITY);
    return f;
}

// Pre-filter blob (spt_filter.cuh): the filterable spheres in index order take consecutive SLOTS; per slot
// (cx, cy, cz, nk) with
//   nk = -(|C|^2 - r^2) + S_i - 1e-3f,  S_i = 96 u (|C|^2 + r^2),  evaluated in binary64 and rounded UP,
// two slots per record, followed by the slot -> sphere index map.  Spheres too large for the filter to say
// anything useful take no slot: they are listed in bigIdx and always evaluated exactly.  Slots are padded to a
// multiple of kQuadSpheres with records that can never pass (index 0xFFFFFFFF).
void build_filter(const SptScene *s, std::vector<unsigned char> &blob, uint32_t &nSlots, std::vector<uint32_t> &big, float &amax)
{
    std::vector<uint32_t> slots;
    big.clear();
    for (uint32_t i = 0; i < s->n; ++i)
    {
        double const cx = s->centers4[4 * i], cy = s->centers4[4 * i + 1], cz = s->centers4[4 * i + 2], r = s->radii[i];
        double const a2 = cx * cx + cy * cy + cz * cz;
        bool const finite = std::isfinite(a2) && std::isfinite(r);
        if (!finite || std::fabs(r) > spt::kBigRadius || std::sqrt(a2) > spt::kBigCentre)
            big.push_back(i);
        else
            slots.push_back(i);
    }
    nSlots = uint32_t((slots.size() + spt::kQuadSpheres - 1) / spt::kQuadSpheres * spt::kQuadSpheres);
    blob.assign(spt::filter_blob_bytes(nSlots), 0);
    spt::FilterPair *pairs = reinterpret_cast<spt::FilterPair *>(blob.data());
    uint32_t *slotIdx = reinterpret_cast<uint32_t *>(pairs + nSlots / 2);
    double amaxD = 0.0;
    for (uint32_t k = 0; k < nSlots; ++k)
    {
        spt::FilterPair &rec = pairs[k / 2];
        int const h = int(k & 1);
        rec.cx[h] = rec.cy[h] = rec.cz[h] = 0.f;
        rec.nk[h] = -1e30f; // never a candidate
        slotIdx[k] = 0xFFFFFFFFu;
        if (k >= slots.size())
            continue;
        uint32_t const i = slots[k];
        double const cx = s->centers4[4 * i], cy = s->centers4[4 * i + 1], cz = s->centers4[4 * i + 2], r = s->radii[i];
        double const a2 = cx * cx + cy * cy + cz * cz;
        double const slack = spt::kSlackU * (a2 + r * r);
        rec.cx[h] = float(cx);
        rec.cy[h] = float(cy);
        rec.cz[h] = float(cz);
        rec.nk[h] = round_up_to_float(-(a2 - r * r) + slack - double(1e-3f));
        slotIdx[k] = i;
        amaxD = std::max(amaxD, std::sqrt(a2));
    }
    amax = round_up_to_float(amaxD);
}

// Bound on |D(xi) - D(0)| over the camera rays of ONE pixel -- jitter xi in [-1, 1)^2, quirk Q2 -- for any pixel of a
// width x height image (spt_device.cuh, primary_direction): r(px, py) = px c0 + py c1 + c2 with px = -1 + 2 (x + xi_v) / H,
// py = -1 + 2 (y + xi_u) / W (swapped divisors, quirk Q1), D = r / |r|.  |r(xi) - r(0)| <= 2 |c0| / H + 2 |c1| / W =: dr, and
// |a/|a| - b/|b|| <= 2 |a - b| / max(|a|, |b|), so delta = 2 dr / min|r| over the image (the exact minimum of a convex
// quadratic over the box of (px, py)), plus the float rounding of the two directions.  4 = "every sphere is a candidate".
double pixel_cone_delta(const SptCamera *cam, uint32_t width, uint32_t height)
{
    double const c0[3] = {cam->view[0], cam->view[4], cam->view[8]}, c1[3] = {cam->view[1], cam->view[5], cam->view[9]},
                 c2[3] = {cam->view[2], cam->view[6], cam->view[10]};
    auto dot = [](const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; };
    double const W = width, H = height;
    double const lo[2] = {-1.0 - 2.0 / H - 1e-6, -1.0 - 2.0 / W - 1e-6}, hi[2] = {-1.0 + 2.0 * W / H + 1e-6, -1.0 + 2.0 * H / W + 1e-6};
    // f(px, py) = |px c0 + py c1 + c2|^2 = a px^2 + 2 b px py + c py^2 + 2 d px + 2 e py + g
    double const a = dot(c0, c0), b = dot(c0, c1), c = dot(c1, c1), d = dot(c0, c2), e = dot(c1, c2), g = dot(c2, c2);
    auto f = [&](double px, double py) { return a * px * px + 2 * b * px * py + c * py * py + 2 * d * px + 2 * e * py + g; };
    auto clamp = [](double v, double l, double h) { return v < l ? l : (v > h ? h : v); };
    double best = std::min(std::min(f(lo[0], lo[1]), f(lo[0], hi[1])), std::min(f(hi[0], lo[1]), f(hi[0], hi[1])));
    double const det = a * c - b * b;
    if (det > 1e-12 * (a * c + 1e-300))
    {
        double const px = (-d * c + e * b) / det, py = (-e * a + d * b) / det; // stationary point
        if (px >= lo[0] && px <= hi[0] && py >= lo[1] && py <= hi[1])
            best = std::min(best, f(px, py));
    }
    for (int k = 0; k < 2; ++k) // the four edges: 1-D minimisers, clamped
    {
        double const pxe = k ? hi[0] : lo[0], pye = k ? hi[1] : lo[1];
        if (c > 0)
            best = std::min(best, f(pxe, clamp(-(b * pxe + e) / c, lo[1], hi[1])));
        if (a > 0)
            best = std::min(best, f(clamp(-(b * pye + d) / a, lo[0], hi[0]), pye));
    }
    double const dr = 2.0 * std::sqrt(a) / H + 2.0 * std::sqrt(c) / W;
    if (!(best > 1e-12) || !std::isfinite(best) || !std::isfinite(dr))
        return 4.0;
    double const rmin = std::sqrt(best) * 0.999;
    if (dr >= 0.25 * rmin)
        return 4.0;
    return 2.0 * dr / rmin * 1.01 + 1e-6;
}

// Camera records of split rendering (spt_filter.cuh, filter_pair_cam: v = tC^2 + nkc with the tC chain of the general form).
// They serve TWO uses in primary_hits_*: once per pixel and launch with the pixel's CENTRAL ray Dc (no jitter), to collect
// the spheres that ANY camera sample of that pixel can hit, and -- for pixels whose list overflows -- per sample as before.
// Per slot, with L = C - E, delta = pixel_cone_delta (0 would give the plain camera form), S = S_i + S_ray(E) the slack of
// the general form (u = 2^-24: covers the exact sequence's and the chain's rounding, DESIGN.md "pre-filter bound"):
//     exact test passes for a sample's ray D   =>   dist(C, line(E, D))^2 < r^2 + 2 S          (2 S: |D| is 1 within 2^-22)
//     dist(C, line(E, Dc)) <= dist(C, line(E, D)) + |L| |D - Dc| < sqrt(r^2 + 2 S) + |L| delta =: R
//     nkc = R^2 - |L|^2 + 2 S   (binary64, rounded up)   =>   v(Dc) = tC(Dc)^2 + nkc >= 0 in the chain's arithmetic.
// S is evaluated with R in place of r.  Conservative for the per-sample use a fortiori (R >= r).
int ensure_camera_blob(SptContext *c, const SptCamera *cam, uint32_t width, uint32_t height)
{
    const float *eye = cam->eye;
    float key[16];
    for (int k = 0; k < 12; ++k)
        key[k] = cam->view[k];
    key[12] = eye[0], key[13] = eye[1], key[14] = eye[2];
    key[15] = 0.f;
    if (c->dPairsCam && c->camW == width && c->camH == height && memcmp(key, c->camKey, sizeof(key)) == 0)
        return SPT_OK;
    uint32_t const nSlots = c->filter.nSlots;
    std::vector<unsigned char> blob = c->hostBlob;
    spt::FilterPair *pairs = reinterpret_cast<spt::FilterPair *>(blob.data());
    double const ex = eye[0], ey = eye[1], ez = eye[2];
    double const oo = ex * ex + ey * ey + ez * ez, B = std::sqrt(oo), m = double(c->filter.amax) + B;
    double sRay = oo * spt::kSlackU * 1.0002;
    double epsT = 0.0;
    if (m > 1024.0)
    {
        epsT = m * 9.5367431640625e-7 * 1.0000002;  // 2^-20 (Amax + B), not below the device's rounded-up value
        sRay += m * m * 3.8146973e-6 * 1.000001;    // + 2^-18 (Amax + B)^2
    }
    double const delta = pixel_cone_delta(cam, width, height);
    for (uint32_t k = 0; k < nSlots; ++k)
    {
        double const cx = c->hostGeom[4 * k], cy = c->hostGeom[4 * k + 1], cz = c->hostGeom[4 * k + 2], r = c->hostGeom[4 * k + 3];
        if (!(r >= 0.0)) // padding slot: stays never-pass
            continue;
        double const a2 = cx * cx + cy * cy + cz * cz;
        double const dx = cx - ex, dy = cy - ey, dz = cz - ez;
        double const l2 = dx * dx + dy * dy + dz * dz, l = std::sqrt(l2);
        double R = r, S = 0.0;
        for (int it = 0; it < 2; ++it) // the slack with the fattened radius in place of r
        {
            S = spt::kSlackU * (a2 + R * R) + sRay;
            R = std::sqrt(r * r + 2.0 * S) + l * delta;
        }
        pairs[k / 2].nk[k & 1] = round_up_to_float((R * R - l2 + 2.0 * S) * (1.0 + 1e-12) + 1e-30);
    }
    // into the arena that is not in use (earlier launches may still read the other one), ordered by the stream
    int const next = c->camArena ^ 1;
    if (blob.size() + 16 > c->camArenaBytes[next])
    {
        SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
        SPT_CUDA_CHECK(cudaStreamSynchronize(c->auxStream));
        unsigned char *fresh = nullptr;
        SPT_CUDA_CHECK(cudaMalloc(&fresh, blob.size() + 16));
        cudaFree(c->dCamArena[next]);
        c->dCamArena[next] = fresh;
        c->camArenaBytes[next] = blob.size() + 16;
    }
    if (!blob.empty())
        SPT_CUDA_CHECK(cudaMemcpyAsync(c->dCamArena[next], blob.data(), blob.size(), cudaMemcpyHostToDevice, c->stream));
    c->camArena = next;
    c->dPairsCam = reinterpret_cast<spt::FilterPair *>(c->dCamArena[next]);
    memcpy(c->camKey, key, sizeof(key));
    c->camW = width;
    c->camH = height;
    c->camEpsT = round_up_to_float(epsT);
    return SPT_OK;
}

// SPT_FLAG_GRID: the uniform grid of spt_grid.cuh over the spheres that have a pre-filter slot (the big ones stay out and are
// always evaluated exactly).  Everything here is in binary64 and rounded towards "more cells per sphere"; the derivation of the
// paddings is in spt_grid.cuh.  gridUsable = false (nothing to put into a grid) makes the flag a no-op.
int ensure_grid(SptContext *c)
{
    if (c->gridValid)
        return SPT_OK;
    c->gridValid = true;
    c->gridUsable = false;
    uint32_t const nSlots = c->filter.nSlots;
    const uint32_t *slotIdx = reinterpret_cast<const uint32_t *>(c->hostBlob.data() + size_t(nSlots) / 2 * sizeof(spt::FilterPair));
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300}, amax = 0.0, rmax = 0.0, reach = 0.0;
    uint32_t count = 0;
    for (uint32_t k = 0; k < nSlots; ++k)
    {
        double const *s = &c->hostGeom[4 * size_t(k)];
        if (!(s[3] >= 0.0) || slotIdx[k] >= c->n)
            continue; // padding
        double const a = std::sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]);
        amax = std::max(amax, a);
        rmax = std::max(rmax, s[3]);
        reach = std::max(reach, a + s[3]);
        for (int d = 0; d < 3; ++d)
        {
            lo[d] = std::min(lo[d], s[d] - s[3]);
            hi[d] = std::max(hi[d], s[d] + s[3]);
        }
        ++count;
    }
    if (count == 0)
        return SPT_OK;
    double const u = 1.0 / 16777216.0, kE = 42.0; // 33.4 u (A^2 + B^2) + 2 u r^2 is the bound; a quarter more
    double const bmax = 2.0 * reach + 16.0; // origins on the listed spheres lie within `reach`; far-out ground contacts do not
    double const margin = std::max(1e-3, (amax + bmax) / 65536.0) * 2.0;
    auto pad_of = [&](double a, double r) { return std::sqrt(r * r + u * (kE * (a * a + bmax * bmax) + 4.0 * r * r)) + margin; };
    double const padMax = pad_of(amax, rmax);
    for (int d = 0; d < 3; ++d)
    {
        lo[d] -= padMax + margin;
        hi[d] += padMax + margin;
    }
    // about two cells per sphere, never finer than the fattest sphere's padding, at most 256 cells per axis
    double const ext[3] = {hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]};
    double cellSize = std::cbrt(ext[0] * ext[1] * ext[2] / (2.0 * count));
    cellSize = std::max(cellSize, std::max(1e-6, 0.5 * padMax));
    uint32_t dim[3];
    double size[3];
    for (int d = 0; d < 3; ++d)
    {
        dim[d] = uint32_t(std::min(256.0, std::max(1.0, std::ceil(ext[d] / cellSize))));
        size[d] = ext[d] / dim[d];
    }
    spt::GridScene &g = c->grid;
    for (int d = 0; d < 3; ++d)
    {
        // the device recomputes cell boundaries as lo + k * size in float: the float box must contain the binary64 one
        g.lo[d] = -round_up_to_float(-lo[d]);
        g.size[d] = round_up_to_float(size[d]);
        g.inv[d] = float(1.0 / double(g.size[d]));
        g.dim[d] = dim[d];
        g.hi[d] = round_up_to_float(double(g.lo[d]) + double(g.size[d]) * dim[d]);
    }
    size_t const nCells = size_t(dim[0]) * dim[1] * dim[2];
    std::vector<uint32_t> &out = c->hostGrid;
    std::vector<uint32_t> counts(nCells + 1, 0);
    auto cell_range = [&](double const *s, int d, uint32_t &c0, uint32_t &c1) {
        double const pad = pad_of(std::sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]), s[3]);
        double const a = (s[d] - pad - double(g.lo[d])) / double(g.size[d]), b = (s[d] + pad - double(g.lo[d])) / double(g.size[d]);
        c0 = uint32_t(std::min<double>(dim[d] - 1, std::max(0.0, std::floor(a))));
        c1 = uint32_t(std::min<double>(dim[d] - 1, std::max(0.0, std::floor(b))));
    };
    for (int pass = 0; pass < 2; ++pass)
    {
        for (uint32_t k = 0; k < nSlots; ++k)
        {
            double const *s = &c->hostGeom[4 * size_t(k)];
            if (!(s[3] >= 0.0) || slotIdx[k] >= c->n)
                continue;
            uint32_t x0, x1, y0, y1, z0, z1;
            cell_range(s, 0, x0, x1);
            cell_range(s, 1, y0, y1);
            cell_range(s, 2, z0, z1);
            for (uint32_t z = z0; z <= z1; ++z)
                for (uint32_t y = y0; y <= y1; ++y)
                    for (uint32_t x = x0; x <= x1; ++x)
                    {
                        size_t const cell = (size_t(z) * dim[1] + y) * dim[0] + x;
                        if (pass == 0)
                            ++counts[cell + 1];
                        else
                            out[nCells + 1 + counts[cell]++] = slotIdx[k];
                    }
        }
        if (pass == 0)
        {
            for (size_t i = 0; i < nCells; ++i)
                counts[i + 1] += counts[i];
            out.assign(nCells + 1 + counts[nCells], 0);
            for (size_t i = 0; i <= nCells; ++i)
                out[i] = counts[i];
        }
    }
    size_t const bytes = out.size() * sizeof(uint32_t);
    if (bytes > c->gridBytes)
    {
        cudaFree(c->dGrid); // (synchronises the device: earlier launches may still walk the old grid)
        c->dGrid = nullptr;
        c->gridBytes = 0;
        SPT_CUDA_CHECK(cudaMalloc(&c->dGrid, bytes));
        c->gridBytes = bytes;
    }
    else
        SPT_CUDA_CHECK(cudaDeviceSynchronize());
    SPT_CUDA_CHECK(cudaMemcpyAsync(c->dGrid, out.data(), bytes, cudaMemcpyHostToDevice, c->stream));
    g.cellStart = reinterpret_cast<const uint32_t *>(c->dGrid);
    g.refs = g.cellStart + nCells + 1;
    g.bmaxSq = float(bmax * bmax * (1.0 - 1e-6));
    g.fatU = round_up_to_float(kE * u);
    g.rmaxSq = round_up_to_float(rmax * rmax * (1.0 + 4.0 * u));
    g.amaxSq = round_up_to_float(amax * amax);
    g.margin = round_up_to_float(margin);
    g.stopAbs = round_up_to_float((bmax + reach) / 1048576.0 + 1e-6);
    c->gridUsable = true;
    return SPT_OK;
}

int check_camera(const SptCamera *cam)
{
    if (!cam)
        return fail(SPT_ERR_INVALID, "camera is null");
    if (cam->view[12] != 0.f || cam->view[13] != 0.f || cam->view[14] != 0.f || cam->view[15] != 0.f || cam->view[3] != 0.f ||
        cam->view[7] != 0.f || cam->view[11] != 0.f || cam->eye[3] != 0.f)
        return fail(SPT_ERR_INVALID, "camera: view row 3 / column 3 and eye.w must be 0 (as Renderer.hpp:321 produces)");
    return SPT_OK;
}

void fill_camera(spt::CameraConst &d, const SptCamera *cam)
{
    for (int r = 0; r < 3; ++r)
        for (int k = 0; k < 4; ++k)
            d.view[4 * r + k] = cam->view[4 * r + k];
    for (int k = 0; k < 3; ++k)
    {
        d.eye[k] = cam->eye[k];
        d.sky[k] = cam->sky[k];
    }
}

int check_segment(uint32_t width, uint32_t height, SptSegment s)
{
    if (width == 0 || height == 0 || uint64_t(width) * height * 3u > 0xFFFFFFFFull)
        return fail(SPT_ERR_INVALID, "image size out of range");
    if (s.xBegin > s.xEnd || s.yBegin > s.yEnd || s.xEnd > width || s.yEnd > height)
        return fail(SPT_ERR_INVALID, "segment outside the image");
    return SPT_OK;
}

int launch_trace(SptContext *c, const SptCamera *cam, const SptParams *prm, SptSegment seg, float4 *dAccum)
{
    if (c->n == 0 || !c->dGeom)
        return fail(SPT_ERR_NO_SCENE, "spt_cuda_set_scene has not been called");
    if (prm->variant != SPT_VARIANT_MEGAKERNEL && prm->variant != SPT_VARIANT_WAVEFRONT)
        return fail(SPT_ERR_INVALID, "unknown variant");

    c->timing.trace_ms = 0.f;
    c->timing.launches = 0;
    c->timing.primary_ms = c->timing.secondary_ms = 0.f;
    c->timedChunks = 0;
    uint32_t const segW = seg.xEnd - seg.xBegin, segH = seg.yEnd - seg.yBegin;
    if (segW == 0 || segH == 0 || prm->sample_end <= prm->sample_begin)
        return SPT_OK;
    // interleaved row bands (SptParams.band_count / band_index): the grid covers only this call's bands
    uint32_t const bandCount = prm->band_count > 1u ? prm->band_count : 1u;
    uint32_t const bandIndex = prm->band_count > 1u ? prm->band_index : 0u;
    if (bandIndex >= bandCount)
        return fail(SPT_ERR_INVALID, "band_index must be below band_count");
    uint32_t const nBands = (segH + SPT_BAND_ROWS - 1u) / SPT_BAND_ROWS;
    uint32_t const myBands = nBands > bandIndex ? (nBands - bandIndex + bandCount - 1u) / bandCount : 0u;
    if (myBands == 0)
        return SPT_OK;

    spt::TraceParams p{};
    p.bandCount = bandCount;
    p.bandIndex = bandIndex;
    fill_camera(p.cam, cam);
    p.width = prm->width;
    p.height = prm->height;
    p.yBegin = seg.yBegin;
    p.yEnd = seg.yEnd;
    p.xBegin = seg.xBegin;
    p.xEnd = seg.xEnd;
    p.sampleBegin = prm->sample_begin;
    p.sampleEnd = prm->sample_end;
    p.depth = prm->depth;
    p.passCap = prm->pass_cap;
    p.seed = prm->seed;
    p.n = c->n;
    p.geom = c->dGeom;
    p.shade = c->dShade;
    p.mat = c->dMat;
    p.filter = c->filter;
    p.accum = dAccum;
    p.counters = (prm->flags & SPT_FLAG_COUNTERS) ? c->dCounters : nullptr;

    static_assert(spt::kTileH == SPT_BAND_ROWS && spt::SplitShape::kTileH == SPT_BAND_ROWS, "a band is one tile row");
    dim3 const grid((segW + spt::kTileW - 1) / spt::kTileW, myBands);
    dim3 const block(spt::kBlockThreads);
    size_t const geomBytes = size_t(c->n) * sizeof(float4);
    size_t const pairBytes = spt::filter_blob_bytes(c->filter.nSlots); // records + slot->index map
    size_t const budget = c->maxSmemOptin > 2048 ? c->maxSmemOptin - 2048 : 0;
    bool const strict = (prm->flags & SPT_FLAG_STRICT_LOOP) != 0;

    SPT_CUDA_CHECK(cudaEventRecord(c->ev[0], c->stream));
#define SPT_LAUNCH(STAGE, STRICT, SMEM)                                                                              \
    do                                                                                                               \
    {                                                                                                                \
        auto kernel = spt::trace_megakernel<STAGE, STRICT>;                                                          \
        SPT_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(SMEM)));        \
        kernel<<<grid, block, (SMEM), c->stream>>>(p);                                                               \
    } while (0)
    if (prm->variant == SPT_VARIANT_WAVEFRONT)
    {
        // persistent CTAs pulling 32x8 tiles from a counter; static shared memory holds the path state
        size_t const wfStatic = 30 * 1024;
        size_t const wfBudget = budget > wfStatic ? budget - wfStatic : 0;
        uint32_t const tiles = grid.x * grid.y;
        uint32_t *tileCounter = reinterpret_cast<uint32_t *>(c->dCounters + 7);
        SPT_CUDA_CHECK(cudaMemsetAsync(tileCounter, 0, sizeof(uint32_t), c->stream));
#define SPT_LAUNCH_WF(STAGE, SMEM)                                                                                    \
    do                                                                                                               \
    {                                                                                                                \
        auto kernel = spt::trace_wavefront<STAGE>;                                                                   \
        SPT_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(SMEM)));        \
        int perSm = 1;                                                                                               \
        SPT_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kernel, spt::kBlockThreads, (SMEM)));   \
        uint32_t const ctas = std::min<uint32_t>(tiles, uint32_t(std::max(perSm, 1)) * uint32_t(c->prop.multiProcessorCount)); \
        kernel<<<ctas, block, (SMEM), c->stream>>>(p, tileCounter);                                                  \
    } while (0)
        if ((prm->flags & SPT_FLAG_STREAMED) && c->filter.nSlots > 0)
            SPT_LAUNCH_WF(3, spt::kStreamSmemBytes);
        else if (geomBytes + pairBytes <= wfBudget)
            SPT_LAUNCH_WF(2, geomBytes + pairBytes);
        else if (pairBytes <= wfBudget)
            SPT_LAUNCH_WF(1, pairBytes);
        else if (c->filter.nSlots > 0)
            SPT_LAUNCH_WF(3, spt::kStreamSmemBytes); // config 5: the blob streams through the chunk ring
        else
            SPT_LAUNCH_WF(0, 0);
#undef SPT_LAUNCH_WF
    }
    else if (strict)
    {
        if (geomBytes <= budget)
            SPT_LAUNCH(1, true, geomBytes);
        else
            SPT_LAUNCH(0, true, 0);
    }
    else
    {
        bool const streamed = (prm->flags & SPT_FLAG_STREAMED) ||
                              (geomBytes + pairBytes > SPT_RESIDENT_LIMIT_BYTES && !(prm->flags & SPT_FLAG_RESIDENT) && c->filter.nSlots > 0);
        bool const fitsResident = geomBytes + pairBytes <= SPT_RESIDENT_LIMIT_BYTES;
        bool const split = !(prm->flags & SPT_FLAG_SINGLE_KERNEL) && (streamed || fitsResident) && c->filter.nSlots > 0;
        size_t const smem = streamed ? size_t(spt::kStreamSmemBytes) : geomBytes + pairBytes;
        if (split)
        {
            // split rendering (spt_split.cuh): per chunk of samples, the primary kernel (camera form of the filter, all lanes
            // in step) writes the first-hit index of every camera sample, then the secondary kernel runs the bounce loop
            // on secondary rays only.  The per-pixel running sums continue across chunks.
            int rc = ensure_camera_blob(c, cam, prm->width, prm->height);
            if (rc)
                return rc;
            p.filterCam = c->filter;
            p.filterCam.pairs = c->dPairsCam;
            p.filterCam.slotIdx = reinterpret_cast<const uint32_t *>(c->dPairsCam + c->filter.nSlots / 2);
            p.filterCam.epsTCam = c->camEpsT;
            // Staging and work items.  The camera kernel stages one 16-byte record per pixel and sample; the budget is the caller's
            // (spt_cuda_set_staging_limit, 8 GiB unless raised), never a share of whatever the device has free.  A call that does
            // not fit is cut into work items = (strip of tile rows) x (chunk of samples); two staging buffers alternate.  Measured
            // on config 3 (profiles/r2_ab_work_items.txt): a lane of the bounce kernel owns its pixel for a whole launch and the
            // warp waits for its slowest lane, so more samples per launch keep more lanes busy (88.7 % of the lanes hold a ray at
            // 32 samples per launch, 94.5 % at 256, 96.0 % at 1024) -- but a launch's CTAs then run proportionally longer, and with
            // the strips that a fixed budget forces (few waves of long CTAs) the tail of every launch costs more than the
            // balance gains: 8 GiB as 270 rows x 32 samples 8 368 Msamples/s, 68 x 128 8 149, 34 x 256 7 852, 17 x 512 7 487.  So
            // the cut is along the SAMPLES, whole image per launch; strips only when not even 16 samples of the image fit.
            // Streams: camera kernels on the first auxiliary stream (the tail of a bounce kernel fills with the next camera
            // kernel's CTAs); strips touch disjoint pixels, so the bounce kernels of consecutive strips run side by side (main
            // and second auxiliary stream); chunks of one strip stay in stream order (the per-pixel running sums continue).
            uint32_t const total = prm->sample_end - prm->sample_begin;
            size_t const rowBytes = size_t(segW) * SPT_BAND_ROWS * sizeof(float4); // one sample of every pixel of one tile row
            size_t const pad = 256;                                                // the consumer prefetches four records past the one it reads
            dim3 const sGridFull = streamed ? grid : dim3((segW + spt::SplitShape::kTileW - 1) / spt::SplitShape::kTileW, myBands);
            uint32_t stripRows = myBands, chunk = total;
            if (prm->split_chunk)
                chunk = std::min(chunk, prm->split_chunk);
            bool const fitsWhole = !prm->strip_rows && size_t(chunk) * myBands * rowBytes + pad <= size_t(c->stagingLimit);
            if (prm->strip_rows)
                stripRows = std::min(myBands, prm->strip_rows);
            else if (!fitsWhole)
            {
                size_t const half = size_t(c->stagingLimit) / 2 > pad ? size_t(c->stagingLimit) / 2 - pad : 0;
                uint32_t const minChunk = std::min(chunk, 16u);
                size_t const whole = half / (size_t(myBands) * rowBytes); // samples of the whole image one buffer holds
                if (whole >= minChunk)
                    chunk = uint32_t(std::min<size_t>(chunk, whole));
                else
                { // a very large image or a very small budget: strips of minChunk samples, shorter chunks only below one tile row
                    chunk = minChunk;
                    stripRows = uint32_t(std::min<size_t>(myBands, half / (size_t(chunk) * rowBytes)));
                    if (stripRows == 0)
                    {
                        stripRows = 1;
                        chunk = uint32_t(std::max<size_t>(1, std::min<size_t>(chunk, half / rowBytes)));
                    }
                }
            }
            uint32_t const nStrips = (myBands + stripRows - 1) / stripRows, nChunks = (total + chunk - 1) / chunk;
            uint32_t const nItems = nStrips * nChunks;
            uint32_t const nBuffers = nItems > 1 ? 2u : 1u;
            size_t const bufBytes = (size_t(chunk) * stripRows * rowBytes + pad + 255) & ~size_t(255);
            if (bufBytes * nBuffers > c->firstHitsBytes)
            {
                SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
                SPT_CUDA_CHECK(cudaStreamSynchronize(c->auxStream));
                SPT_CUDA_CHECK(cudaStreamSynchronize(c->auxStream2));
                cudaFree(c->dFirstHits);
                c->dFirstHits = nullptr;
                c->firstHitsBytes = 0;
                cudaError_t const ce = cudaMalloc(&c->dFirstHits, bufBytes * nBuffers);
                if (ce != cudaSuccess)
                {
                    c->dFirstHits = nullptr;
                    (void)cudaGetLastError();
                    return fail(SPT_ERR_CUDA, "cudaMalloc of the first-hit staging buffers failed (lower spt_cuda_set_staging_limit)", ce);
                }
                c->firstHitsBytes = bufBytes * nBuffers;
            }
            void (*primary)(const spt::TraceParams) = streamed ? spt::primary_hits_streamed : spt::primary_hits_kernel<2, spt::SplitShape>;
            // samples shared inside the warp (spt_split.cuh, AccDeposit) unless the caller opts out or a chunk does not fit the task word
            bool const share = !(prm->flags & SPT_FLAG_OWN_PIXELS) && chunk < spt::kTaskSampleMask;
            void (*secondary)(const spt::TraceParams) =
                streamed ? (share ? spt::trace_secondary_streamed<true> : spt::trace_secondary_streamed<false>)
                         : (share ? spt::trace_secondary<2, spt::SplitShape, true> : spt::trace_secondary<2, spt::SplitShape, false>);
            // SPT_FLAG_GRID (opt-in): the bounce loop's closest hit walks the uniform grid instead of the list
            void (*secondaryGrid)(const spt::TraceParams, const spt::GridScene) = nullptr;
            void (*primaryGrid)(const spt::TraceParams, const spt::GridScene) = nullptr;
            if (prm->flags & SPT_FLAG_GRID)
            {
                rc = ensure_grid(c);
                if (rc)
                    return rc;
                if (c->gridUsable)
                {
                    // The camera kernel walks the grid as well when the list is a long one (on resident lists the pixel-cone
                    // kernel is the faster camera kernel: config 3 20 ms against 69 ms, config 4 8 ms against 17 ms; config 5 78 ms
                    // against 33 ms, session 26) and the eye is within the grid's reach (a far eye would send every ray that points
                    // at the scene through the whole list).
                    double const ee = double(cam->eye[0]) * cam->eye[0] + double(cam->eye[1]) * cam->eye[1] + double(cam->eye[2]) * cam->eye[2];
                    if (streamed && ee <= 0.99 * double(c->grid.bmaxSq))
                        primaryGrid = spt::primary_hits_grid<spt::DefaultShape>;
                }
                if (c->gridUsable)
                    secondaryGrid = streamed ? (share ? spt::trace_secondary_grid<spt::DefaultShape, true> : spt::trace_secondary_grid<spt::DefaultShape, false>)
                                             : (share ? spt::trace_secondary_grid<spt::SplitShape, true> : spt::trace_secondary_grid<spt::SplitShape, false>);
            }
            // the resident kernels run smaller CTAs than the rest (spt_split.cuh, SplitShape)
            dim3 const sBlock(streamed ? spt::kBlockThreads : spt::SplitShape::kThreads);
            SPT_CUDA_CHECK(cudaFuncSetAttribute(primary, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            SPT_CUDA_CHECK(cudaFuncSetAttribute(secondary, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            // P(i) needs buffer i % 2 free (S(i-2) done), S(i) needs P(i); everything the auxiliary streams do is waited for on
            // the main stream before this function returns to its caller's stream order.
            bool const overlap = nBuffers > 1 && !(prm->flags & SPT_FLAG_SERIAL_CHUNKS);
            cudaStream_t const sp = overlap ? c->auxStream : c->stream;
            cudaStream_t const sAlt = (overlap && nStrips > 1) ? c->auxStream2 : c->stream;
            while (c->evKernel.size() < size_t(nItems) * 4)
            {
                cudaEvent_t e = nullptr;
                SPT_CUDA_CHECK(cudaEventCreate(&e));
                c->evKernel.push_back(e);
            }
            c->timedChunks = nItems;
            if (overlap)
            {
                SPT_CUDA_CHECK(cudaEventRecord(c->evFork, c->stream));
                SPT_CUDA_CHECK(cudaStreamWaitEvent(sp, c->evFork, 0));
                if (sAlt != c->stream)
                    SPT_CUDA_CHECK(cudaStreamWaitEvent(sAlt, c->evFork, 0));
            }
            uint32_t i = 0;
            for (uint32_t strip = 0; strip < nStrips; ++strip)
            {
                uint32_t const r0 = strip * stripRows, r1 = std::min(myBands, r0 + stripRows);
                dim3 const sGrid(sGridFull.x, r1 - r0);
                cudaStream_t const ss = (strip & 1u) ? sAlt : c->stream;
                p.tileRowOffset = r0;
                for (uint32_t s0 = prm->sample_begin; s0 < prm->sample_end; s0 += chunk, ++i)
                {
                    uint32_t const b = i % nBuffers;
                    p.sampleBegin = s0;
                    p.sampleEnd = std::min(prm->sample_end, s0 + chunk);
                    p.firstHits = reinterpret_cast<float4 *>(reinterpret_cast<unsigned char *>(c->dFirstHits) + size_t(b) * bufBytes);
                    if (overlap && i >= 2)
                        SPT_CUDA_CHECK(cudaStreamWaitEvent(sp, c->evSecondary[b], 0));
                    SPT_CUDA_CHECK(cudaEventRecord(c->evKernel[4 * i + 0], sp));
                    if (primaryGrid)
                        primaryGrid<<<sGrid, sBlock, 0, sp>>>(p, c->grid);
                    else
                        primary<<<sGrid, sBlock, smem, sp>>>(p);
                    SPT_CUDA_CHECK(cudaEventRecord(c->evKernel[4 * i + 1], sp));
                    if (overlap)
                    {
                        SPT_CUDA_CHECK(cudaEventRecord(c->evPrimary[b], sp));
                        SPT_CUDA_CHECK(cudaStreamWaitEvent(ss, c->evPrimary[b], 0));
                    }
                    SPT_CUDA_CHECK(cudaEventRecord(c->evKernel[4 * i + 2], ss));
                    if (secondaryGrid)
                        secondaryGrid<<<sGrid, sBlock, 0, ss>>>(p, c->grid);
                    else
                        secondary<<<sGrid, sBlock, smem, ss>>>(p);
                    SPT_CUDA_CHECK(cudaEventRecord(c->evKernel[4 * i + 3], ss));
                    if (overlap)
                        SPT_CUDA_CHECK(cudaEventRecord(c->evSecondary[b], ss));
                    c->timing.launches += 2;
                }
            }
            p.tileRowOffset = 0;
            if (sAlt != c->stream)
            {
                SPT_CUDA_CHECK(cudaEventRecord(c->evJoin, sAlt));
                SPT_CUDA_CHECK(cudaStreamWaitEvent(c->stream, c->evJoin, 0));
            }
            c->timing.launches -= 1; // the common tail below counts one
        }
        else if (streamed)
        {
            // list too long to sit in shared memory next to four resident CTAs: TMA chunk ring (spt_streamed.cuh)
            SPT_CUDA_CHECK(cudaFuncSetAttribute(spt::trace_streamed, cudaFuncAttributeMaxDynamicSharedMemorySize, int(spt::kStreamSmemBytes)));
            spt::trace_streamed<<<grid, block, spt::kStreamSmemBytes, c->stream>>>(p);
        }
        else if (geomBytes + pairBytes <= budget)
            SPT_LAUNCH(2, false, geomBytes + pairBytes);
        else if (pairBytes <= budget)
            SPT_LAUNCH(1, false, pairBytes);
        else
            SPT_LAUNCH(0, false, 0);
    }
#undef SPT_LAUNCH
    SPT_CUDA_CHECK(cudaGetLastError());
    SPT_CUDA_CHECK(cudaEventRecord(c->ev[1], c->stream));
    c->timing.launches += 1;
    return SPT_OK;
}

int launch_resolve(SptContext *c, const float4 *dAccum, uint32_t width, uint32_t height, SptSegment seg, uint8_t *dRgb)
{
    uint32_t const segW = seg.xEnd - seg.xBegin, segH = seg.yEnd - seg.yBegin;
    c->timing.resolve_ms = 0.f;
    if (segW == 0 || segH == 0)
        return SPT_OK;
    dim3 const block(128);
    dim3 const grid((segW + block.x - 1) / block.x, segH);
    SPT_CUDA_CHECK(cudaEventRecord(c->ev[2], c->stream));
    spt::resolve_kernel<<<grid, block, 0, c->stream>>>(dAccum, width, height, seg.yBegin, seg.yEnd, seg.xBegin, seg.xEnd, dRgb);
    SPT_CUDA_CHECK(cudaGetLastError());
    SPT_CUDA_CHECK(cudaEventRecord(c->ev[3], c->stream));
    c->timing.launches += 1;
    return SPT_OK;
}

int create_one(SptContext **out, int device)
{
    int count = 0;
    cudaError_t ce = cudaGetDeviceCount(&count);
    if (ce != cudaSuccess || count <= 0)
        return fail(SPT_ERR_NO_DEVICE, "no CUDA device is usable; this library has no CPU path", ce);
    if (device < 0 || device >= count)
        return fail(SPT_ERR_INVALID, "device ordinal out of range");
    SPT_CUDA_CHECK(cudaSetDevice(device));
    SptContext *c = new SptContext();
    c->device = device;
    ce = cudaGetDeviceProperties(&c->prop, device);
    if (ce == cudaSuccess)
        ce = cudaStreamCreateWithFlags(&c->ownStream, cudaStreamNonBlocking);
    if (ce == cudaSuccess)
        ce = cudaStreamCreateWithFlags(&c->auxStream, cudaStreamNonBlocking);
    if (ce == cudaSuccess)
        ce = cudaStreamCreateWithFlags(&c->auxStream2, cudaStreamNonBlocking);
    for (int i = 0; i < 4 && ce == cudaSuccess; ++i)
        ce = cudaEventCreate(&c->ev[i]);
    cudaEvent_t *const sync[] = {&c->evFork, &c->evJoin, &c->evPrimary[0], &c->evPrimary[1], &c->evSecondary[0], &c->evSecondary[1], &c->evDone};
    for (cudaEvent_t *e : sync)
        if (ce == cudaSuccess)
            ce = cudaEventCreateWithFlags(e, cudaEventDisableTiming);
    if (ce == cudaSuccess)
        ce = cudaMalloc(&c->dCounters, 8 * sizeof(unsigned long long));
    if (ce == cudaSuccess)
        ce = cudaMemset(c->dCounters, 0, 8 * sizeof(unsigned long long));
    if (ce != cudaSuccess)
    {
        delete c;
        return fail(SPT_ERR_CUDA, "context creation", ce);
    }
    c->stream = c->ownStream;
    c->maxSmemOptin = c->prop.sharedMemPerBlockOptin;
    // the static scratch array's address on THIS device (cudaSetDevice(device) above is still in force); a failure is kept
    // for spt_cuda_closest_hit_static to report and does not fail the context
    c->staticScratchError = cudaGetSymbolAddress(&c->dStaticScratch, g_staticScratch);
    if (c->staticScratchError != cudaSuccess)
    {
        c->dStaticScratch = nullptr;
        (void)cudaGetLastError();
    }
    *out = c;
    return SPT_OK;
}

void destroy_one(SptContext *c)
{
    if (!c)
        return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->auxStream)
        cudaStreamSynchronize(c->auxStream);
    cudaFree(c->dSceneArena[0]);
    cudaFree(c->dSceneArena[1]);
    cudaFree(c->dCounters);
    cudaFree(c->dAccum);
    cudaFree(c->dRgb);
    cudaFree(c->dProbe);
    cudaFree(c->dProgAccum);
    cudaFree(c->dProgRgb);
    cudaFree(c->dCamArena[0]);
    cudaFree(c->dCamArena[1]);
    cudaFree(c->dFirstHits);
    cudaFree(c->dGrid);
    for (auto &e : c->ev)
        if (e)
            cudaEventDestroy(e);
    for (cudaEvent_t e : c->evKernel)
        cudaEventDestroy(e);
    for (cudaEvent_t e : {c->evFork, c->evJoin, c->evPrimary[0], c->evPrimary[1], c->evSecondary[0], c->evSecondary[1], c->evDone})
        if (e)
            cudaEventDestroy(e);
    if (c->ownStream)
        cudaStreamDestroy(c->ownStream);
    if (c->auxStream)
        cudaStreamDestroy(c->auxStream);
    if (c->auxStream2)
        cudaStreamDestroy(c->auxStream2);
    delete c;
}

// the scene as every device receives it, prepared once on the host
struct HostScene
{
    uint32_t n = 0;
    std::vector<float4> geom, shade;
    const uint8_t *materials = nullptr;
    std::vector<unsigned char> blob;
    std::vector<uint32_t> big;
    std::vector<double> slotGeom;
    uint32_t nSlots = 0;
    float amax = 0.f;
};

// The scene of a device lives in ONE allocation, packed [geometry | shading | materials | pre-filter blob | big-sphere list],
// and every context keeps TWO such arenas that take turns: a new scene is copied (one asynchronous copy on the context's
// stream) into the arena that is NOT in use, then the pointers switch.  Launches already queued keep reading the old arena;
// in steady state (re-uploading scenes of the same size, as an end-to-end render loop does) nothing is allocated or freed;
// and a failed allocation or copy leaves the context exactly as it was.
int upload_scene(SptContext *c, const HostScene &hs)
{
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    auto align = [](size_t v) { return (v + 255) & ~size_t(255); };
    size_t const oGeom = 0, oShade = align(oGeom + hs.n * sizeof(float4)), oMat = align(oShade + hs.n * sizeof(float4)),
                 oPairs = align(oMat + hs.n), oBig = align(oPairs + hs.blob.size() + 16), total = align(oBig + (hs.big.size() + 1) * sizeof(uint32_t));
    int const next = c->sceneArena ^ 1;
    if (total > c->sceneArenaBytes[next])
    {
        // the arena about to be replaced may still be read by launches queued before the PREVIOUS upload
        SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
        SPT_CUDA_CHECK(cudaStreamSynchronize(c->auxStream));
        unsigned char *fresh = nullptr;
        cudaError_t const ce = cudaMalloc(&fresh, total);
        if (ce != cudaSuccess)
        {
            (void)cudaGetLastError();
            return fail(SPT_ERR_CUDA, "spt_cuda_set_scene: allocation failed, the previous scene is kept", ce);
        }
        cudaFree(c->dSceneArena[next]);
        c->dSceneArena[next] = fresh;
        c->sceneArenaBytes[next] = total;
    }
    // pack on the host, one copy
    c->hostScenePacked.assign(total, 0);
    unsigned char *h = c->hostScenePacked.data();
    memcpy(h + oGeom, hs.geom.data(), hs.n * sizeof(float4));
    memcpy(h + oShade, hs.shade.data(), hs.n * sizeof(float4));
    memcpy(h + oMat, hs.materials, hs.n);
    if (!hs.blob.empty())
        memcpy(h + oPairs, hs.blob.data(), hs.blob.size());
    if (!hs.big.empty())
        memcpy(h + oBig, hs.big.data(), hs.big.size() * sizeof(uint32_t));
    unsigned char *d = c->dSceneArena[next];
    // (pageable source: the call returns once the bytes sit in the driver's staging memory, so `h` may be reused at once)
    cudaError_t const ce = cudaMemcpyAsync(d, h, total, cudaMemcpyHostToDevice, c->stream);
    if (ce != cudaSuccess)
        return fail(SPT_ERR_CUDA, "spt_cuda_set_scene: upload failed, the previous scene is kept", ce);
    c->sceneArena = next;
    c->dGeom = reinterpret_cast<float4 *>(d + oGeom);
    c->dShade = reinterpret_cast<float4 *>(d + oShade);
    c->dMat = d + oMat;
    c->dPairs = reinterpret_cast<spt::FilterPair *>(d + oPairs);
    c->dBigIdx = reinterpret_cast<uint32_t *>(d + oBig);
    c->camKey[0] = NAN; // the camera records are derived from the scene: rebuilt by the next split render
    c->hostBlob = hs.blob;
    c->hostGeom = hs.slotGeom;
    c->gridValid = false;
    c->filter.pairs = c->dPairs;
    c->filter.nSlots = hs.nSlots;
    c->filter.shape = spt::span_shape(hs.nSlots);
    c->filter.slotIdx = reinterpret_cast<const uint32_t *>(c->dPairs + hs.nSlots / 2);
    c->filter.bigIdx = c->dBigIdx;
    c->filter.nBig = uint32_t(hs.big.size());
    c->filter.amax = hs.amax;
    {
        // origins with |O|^2 <= nearOO satisfy amax + |O| <= 1024 (make_filter_ray's tC test needs no widening there)
        double const t = 1024.0 - double(hs.amax);
        float near = -1.f;
        if (t > 0.0)
        {
            near = float(t * t);
            if (double(near) > t * t)
                near = std::nextafterf(near, -INFINITY);
        }
        c->filter.nearOO = near;
    }
    c->n = hs.n;
    return SPT_OK;
}

// segment of the device image / accumulator -> the caller's host buffers (pixel (x,y) lives in byte row H-1-y)
int copy_segment_to_host(SptContext *c, uint32_t W, uint32_t H, SptSegment seg, const uint8_t *dRgb, uint8_t *rgb8, const float4 *dAccum,
                         float *accum4)
{
    uint32_t const segW = seg.xEnd - seg.xBegin, segH = seg.yEnd - seg.yBegin;
    if (segW == 0 || segH == 0)
        return SPT_OK;
    uint32_t const rowTop = H - seg.yEnd; // flipped
    if (rgb8)
        SPT_CUDA_CHECK(cudaMemcpy2DAsync(rgb8 + (size_t(rowTop) * W + seg.xBegin) * 3, size_t(W) * 3,
                                         dRgb + (size_t(rowTop) * W + seg.xBegin) * 3, size_t(W) * 3, size_t(segW) * 3, segH,
                                         cudaMemcpyDeviceToHost, c->stream));
    if (accum4)
        SPT_CUDA_CHECK(cudaMemcpy2DAsync(accum4 + (size_t(seg.yBegin) * W + seg.xBegin) * 4, size_t(W) * sizeof(float4),
                                         dAccum + size_t(seg.yBegin) * W + seg.xBegin, size_t(W) * sizeof(float4),
                                         size_t(segW) * sizeof(float4), segH, cudaMemcpyDeviceToHost, c->stream));
    return SPT_OK;
}

// One call spread over the devices of a multi-device context (Renderer.hpp:257-302's worker pool): device k traces samples
// [b + k S / n, b + (k + 1) S / n) of every pixel of the segment, or -- fewer samples than devices, or by request -- all
// samples of the interleaved 8-row bands b with b % n == k, into accOf(device).  Asynchronous on each device's stream.
template <typename AccOf>
int trace_multi(SptContext *c, const SptCamera *cam, const SptParams *prm, SptSegment seg, AccOf accOf)
{
    if (prm->band_count > 1u)
        return fail(SPT_ERR_INVALID, "band_count / band_index belong to single-device contexts: a multi-device context partitions itself");
    uint32_t const n = uint32_t(c->all.size());
    uint32_t const total = prm->sample_end > prm->sample_begin ? prm->sample_end - prm->sample_begin : 0u;
    bool const bySamples = c->partition == SPT_PARTITION_SAMPLES || (c->partition == SPT_PARTITION_AUTO && total >= n);
    for (uint32_t k = 0; k < n; ++k)
    {
        SptContext *d = c->all[k];
        SPT_CUDA_CHECK(cudaSetDevice(d->device));
        if (!accOf(d))
            return fail(SPT_ERR_INVALID, "accumulator missing on a device of the context");
        SptParams pk = *prm;
        if (bySamples)
        {
            pk.sample_begin = prm->sample_begin + uint32_t(uint64_t(total) * k / n);
            pk.sample_end = prm->sample_begin + uint32_t(uint64_t(total) * (k + 1) / n);
        }
        else
        {
            pk.band_count = n;
            pk.band_index = k;
        }
        int const rc = launch_trace(d, cam, &pk, seg, accOf(d));
        if (rc)
            return rc;
    }
    return SPT_OK;
}

// The exchange step: every device sums one slice of the segment's rows over all devices' accumulators (peer loads over
// NVLink, device order), encodes it, and copies its slice to the host buffers.  Blocking.
template <typename SrcOf, typename SumOf, typename RgbOf>
int reduce_multi(SptContext *c, uint32_t W, uint32_t H, SptSegment seg, SrcOf srcOf, SumOf sumOf, RgbOf rgbOf, uint8_t *rgb8, float *accum4)
{
    uint32_t const n = uint32_t(c->all.size());
    uint32_t const segW = seg.xEnd - seg.xBegin, segH = seg.yEnd - seg.yBegin;
    spt::ReduceParams r{};
    r.n = n;
    for (uint32_t k = 0; k < n; ++k)
    {
        SptContext *d = c->all[k];
        r.acc[k] = srcOf(d);
        SPT_CUDA_CHECK(cudaSetDevice(d->device));
        SPT_CUDA_CHECK(cudaEventRecord(d->evDone, d->stream));
    }
    for (uint32_t k = 0; k < n; ++k)
    {
        SptContext *d = c->all[k];
        SPT_CUDA_CHECK(cudaSetDevice(d->device));
        d->timing.resolve_ms = 0.f;
        uint32_t const y0 = seg.yBegin + uint32_t(uint64_t(segH) * k / n), y1 = seg.yBegin + uint32_t(uint64_t(segH) * (k + 1) / n);
        if (y1 == y0)
            continue;
        for (uint32_t j = 0; j < n; ++j)
            if (j != k)
                SPT_CUDA_CHECK(cudaStreamWaitEvent(d->stream, c->all[j]->evDone, 0));
        dim3 const block(128);
        dim3 const grid((segW + block.x - 1) / block.x, y1 - y0);
        SPT_CUDA_CHECK(cudaEventRecord(d->ev[2], d->stream));
        spt::reduce_resolve_kernel<<<grid, block, 0, d->stream>>>(r, W, H, y0, y1, seg.xBegin, seg.xEnd, accum4 ? sumOf(d) : nullptr,
                                                                   rgb8 ? rgbOf(d) : nullptr);
        SPT_CUDA_CHECK(cudaGetLastError());
        SPT_CUDA_CHECK(cudaEventRecord(d->ev[3], d->stream));
        d->timing.launches += 1;
        int const rc = copy_segment_to_host(d, W, H, SptSegment{y0, y1, seg.xBegin, seg.xEnd}, rgbOf(d), rgb8, sumOf(d), accum4);
        if (rc)
            return rc;
    }
    for (SptContext *d : c->all)
    {
        SPT_CUDA_CHECK(cudaSetDevice(d->device));
        SPT_CUDA_CHECK(cudaStreamSynchronize(d->stream));
    }
    return SPT_OK;
}
} // namespace

extern "C"
{

const char *spt_cuda_last_error(void) { return g_lastError.c_str(); }
const char *spt_cuda_version(void) { return "simplepathtracer_b200 0.1 (sm_100a)"; }
uint32_t spt_cuda_abi_version(void) { return SPT_ABI_VERSION; }

int spt_cuda_create(SptContext **out, int device)
{
    if (!out)
        return fail(SPT_ERR_INVALID, "out is null");
    *out = nullptr;
    SptContext *c = nullptr;
    int const rc = create_one(&c, device);
    if (rc)
        return rc;
    c->all.assign(1, c);
    *out = c;
    return SPT_OK;
}

int spt_cuda_create_multi(SptContext **out, const int *devices, int n_devices)
{
    if (!out || !devices)
        return fail(SPT_ERR_INVALID, "null argument");
    *out = nullptr;
    if (n_devices < 1 || n_devices > SPT_MAX_DEVICES)
        return fail(SPT_ERR_INVALID, "n_devices must be in 1..SPT_MAX_DEVICES");
    static_assert(SPT_MAX_DEVICES <= spt::kMaxDevices, "ReduceParams holds one pointer per device");
    std::vector<SptContext *> all;
    auto cleanup = [&]() {
        for (SptContext *d : all)
            destroy_one(d);
    };
    for (int k = 0; k < n_devices; ++k)
    {
        SptContext *d = nullptr;
        int const rc = create_one(&d, devices[k]);
        if (rc)
        {
            cleanup();
            return rc;
        }
        all.push_back(d);
    }
    // every device reads every other device's accumulator in the fused reduce-and-resolve kernel
    for (SptContext *a : all)
        for (SptContext *b : all)
        {
            if (a->device == b->device)
                continue;
            int can = 0;
            cudaError_t ce = cudaDeviceCanAccessPeer(&can, a->device, b->device);
            if (ce == cudaSuccess && !can)
            {
                cleanup();
                return fail(SPT_ERR_CUDA, "spt_cuda_create_multi: the devices have no peer access to each other");
            }
            if (ce == cudaSuccess)
                ce = cudaSetDevice(a->device);
            if (ce == cudaSuccess)
            {
                ce = cudaDeviceEnablePeerAccess(b->device, 0);
                if (ce == cudaErrorPeerAccessAlreadyEnabled)
                {
                    (void)cudaGetLastError();
                    ce = cudaSuccess;
                }
            }
            if (ce != cudaSuccess)
            {
                cleanup();
                return fail(SPT_ERR_CUDA, "cudaDeviceEnablePeerAccess", ce);
            }
        }
    all[0]->all = all;
    *out = all[0];
    return SPT_OK;
}

int spt_cuda_device_count(SptContext *c) { return c ? int(c->all.size()) : 0; }

int spt_cuda_set_partition(SptContext *c, int partition)
{
    if (!c)
        return fail(SPT_ERR_INVALID, "context is null");
    if (partition != SPT_PARTITION_AUTO && partition != SPT_PARTITION_SAMPLES && partition != SPT_PARTITION_BANDS)
        return fail(SPT_ERR_INVALID, "unknown partition");
    std::lock_guard<std::mutex> g(c->lock);
    c->partition = partition;
    return SPT_OK;
}

int spt_cuda_set_staging_limit(SptContext *c, uint64_t bytes)
{
    if (!c)
        return fail(SPT_ERR_INVALID, "context is null");
    std::lock_guard<std::mutex> g(c->lock);
    for (SptContext *d : c->all)
        d->stagingLimit = bytes ? bytes : SPT_DEFAULT_STAGING_BYTES;
    return SPT_OK;
}

void spt_cuda_destroy(SptContext *c)
{
    if (!c)
        return;
    std::vector<SptContext *> const all = c->all;
    for (SptContext *d : all)
        destroy_one(d);
}

int spt_cuda_set_stream(SptContext *c, void *cuda_stream, int use_own)
{
    if (!c)
        return fail(SPT_ERR_INVALID, "context is null");
    std::lock_guard<std::mutex> g(c->lock);
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    // work queued on the stream being left may still use context-owned buffers (staging, camera records, scratch)
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->auxStream));
    // a null handle is a real stream (the legacy default stream), so "own" is a separate switch
    c->stream = use_own ? c->ownStream : static_cast<cudaStream_t>(cuda_stream);
    return SPT_OK;
}

int spt_cuda_device_info(SptContext *c, SptDeviceInfo *out)
{
    if (!c || !out)
        return fail(SPT_ERR_INVALID, "null argument");
    memset(out, 0, sizeof(*out));
    strncpy(out->name, c->prop.name, sizeof(out->name) - 1);
    out->sm_count = c->prop.multiProcessorCount;
    out->cc_major = c->prop.major;
    out->cc_minor = c->prop.minor;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, c->device);
    out->clock_khz = khz;
    out->smem_per_block_optin = c->prop.sharedMemPerBlockOptin;
    out->global_mem = c->prop.totalGlobalMem;
    return SPT_OK;
}

int spt_cuda_set_scene(SptContext *c, const SptScene *s)
{
    if (!c || !s)
        return fail(SPT_ERR_INVALID, "null argument");
    if (s->n == 0 || !s->centers4 || !s->radii || !s->materials || !s->colors4 || !s->fuzz)
        return fail(SPT_ERR_INVALID, "scene arrays must be non-null and n > 0");
    std::lock_guard<std::mutex> g(c->lock);
    HostScene hs;
    hs.n = s->n;
    hs.materials = s->materials;
    hs.geom.resize(s->n);
    hs.shade.resize(s->n);
    for (uint32_t i = 0; i < s->n; ++i)
    {
        if (s->centers4[4 * i + 3] != 0.f)
            return fail(SPT_ERR_INVALID, "sphere centre w must be 0 (Vec4{x,y,z} initialisers, SceneGenerators.hpp)");
        hs.geom[i] = make_float4(s->centers4[4 * i], s->centers4[4 * i + 1], s->centers4[4 * i + 2], s->radii[i]);
        hs.shade[i] = make_float4(s->colors4[4 * i], s->colors4[4 * i + 1], s->colors4[4 * i + 2], s->fuzz[i]);
    }
    build_filter(s, hs.blob, hs.nSlots, hs.big, hs.amax);
    // host copies from which the camera form of the records is derived per eye (ensure_camera_blob)
    hs.slotGeom.assign(size_t(hs.nSlots) * 4, -1.0);
    {
        const uint32_t *slotIdx = reinterpret_cast<const uint32_t *>(hs.blob.data() + size_t(hs.nSlots) / 2 * sizeof(spt::FilterPair));
        for (uint32_t k = 0; k < hs.nSlots; ++k)
        {
            uint32_t const i = slotIdx[k];
            if (i >= s->n)
                continue; // padding: radius stays -1
            hs.slotGeom[4 * k + 0] = s->centers4[4 * i];
            hs.slotGeom[4 * k + 1] = s->centers4[4 * i + 1];
            hs.slotGeom[4 * k + 2] = s->centers4[4 * i + 2];
            hs.slotGeom[4 * k + 3] = std::fabs(double(s->radii[i]));
        }
    }
    for (SptContext *d : c->all)
    {
        int const rc = upload_scene(d, hs);
        if (rc)
            return rc;
    }
    return SPT_OK;
}

int spt_cuda_accumulate(SptContext *c, const SptCamera *cam, const SptParams *prm, SptSegment seg, float *d_accum4)
{
    if (!c || !prm || !d_accum4)
        return fail(SPT_ERR_INVALID, "null argument");
    int rc = check_camera(cam);
    if (rc)
        return rc;
    rc = check_segment(prm->width, prm->height, seg);
    if (rc)
        return rc;
    std::lock_guard<std::mutex> g(c->lock);
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    return launch_trace(c, cam, prm, seg, reinterpret_cast<float4 *>(d_accum4));
}

int spt_cuda_resolve(SptContext *c, const float *d_accum4, uint32_t width, uint32_t height, SptSegment seg, uint8_t *d_rgb8)
{
    if (!c || !d_accum4 || !d_rgb8)
        return fail(SPT_ERR_INVALID, "null argument");
    int const rc = check_segment(width, height, seg);
    if (rc)
        return rc;
    std::lock_guard<std::mutex> g(c->lock);
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    return launch_resolve(c, reinterpret_cast<const float4 *>(d_accum4), width, height, seg, d_rgb8);
}

int spt_cuda_synchronize(SptContext *c)
{
    if (!c)
        return fail(SPT_ERR_INVALID, "context is null");
    std::lock_guard<std::mutex> g(c->lock);
    for (SptContext *d : c->all)
    {
        SPT_CUDA_CHECK(cudaSetDevice(d->device));
        SPT_CUDA_CHECK(cudaStreamSynchronize(d->stream));
    }
    return SPT_OK;
}

int spt_cuda_render(SptContext *c, const SptCamera *cam, const SptParams *prm, SptSegment seg, uint8_t *rgb8, float *accum4)
{
    if (!c || !prm || !rgb8)
        return fail(SPT_ERR_INVALID, "null argument");
    int rc = check_camera(cam);
    if (rc)
        return rc;
    rc = check_segment(prm->width, prm->height, seg);
    if (rc)
        return rc;
    std::lock_guard<std::mutex> g(c->lock);
    uint32_t const W = prm->width, H = prm->height;
    uint32_t const segW = seg.xEnd - seg.xBegin, segH = seg.yEnd - seg.yBegin;
    if (segW == 0 || segH == 0)
        return SPT_OK;
    if (c->all.size() > 1)
    {
        // every device: zero its accumulator rows, trace its share
        for (SptContext *d : c->all)
        {
            SPT_CUDA_CHECK(cudaSetDevice(d->device));
            rc = ensure_scratch(d, size_t(W) * H);
            if (rc)
                return rc;
            SPT_CUDA_CHECK(cudaMemset2DAsync(d->dAccum + size_t(seg.yBegin) * W + seg.xBegin, size_t(W) * sizeof(float4), 0,
                                             size_t(segW) * sizeof(float4), segH, d->stream));
        }
        rc = trace_multi(c, cam, prm, seg, [](SptContext *d) { return d->dAccum; });
        if (rc)
            return rc;
        // the totals may replace each device's own slice of its accumulator
        return reduce_multi(c, W, H, seg, [](SptContext *d) { return d->dAccum; }, [](SptContext *d) { return d->dAccum; },
                            [](SptContext *d) { return d->dRgb; }, rgb8, accum4);
    }
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    rc = ensure_scratch(c, size_t(W) * H);
    if (rc)
        return rc;
    // zero the accumulator rows of the segment
    SPT_CUDA_CHECK(cudaMemset2DAsync(c->dAccum + size_t(seg.yBegin) * W + seg.xBegin, size_t(W) * sizeof(float4), 0,
                                     size_t(segW) * sizeof(float4), segH, c->stream));
    rc = launch_trace(c, cam, prm, seg, c->dAccum);
    if (rc)
        return rc;
    rc = launch_resolve(c, c->dAccum, W, H, seg, c->dRgb);
    if (rc)
        return rc;
    rc = copy_segment_to_host(c, W, H, seg, c->dRgb, rgb8, c->dAccum, accum4);
    if (rc)
        return rc;
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    return SPT_OK;
}

int spt_cuda_progressive_begin(SptContext *c, uint32_t width, uint32_t height)
{
    if (!c)
        return fail(SPT_ERR_INVALID, "context is null");
    int const rc = check_segment(width, height, SptSegment{0, height, 0, width});
    if (rc)
        return rc;
    std::lock_guard<std::mutex> g(c->lock);
    for (SptContext *d : c->all)
    {
        SPT_CUDA_CHECK(cudaSetDevice(d->device));
        SPT_CUDA_CHECK(cudaStreamSynchronize(d->stream));
        if (width != d->progW || height != d->progH || !d->dProgAccum)
        {
            cudaFree(d->dProgAccum);
            cudaFree(d->dProgRgb);
            d->dProgAccum = nullptr;
            d->dProgRgb = nullptr;
            d->progW = d->progH = 0;
            SPT_CUDA_CHECK(cudaMalloc(&d->dProgAccum, size_t(width) * height * sizeof(float4)));
            SPT_CUDA_CHECK(cudaMalloc(&d->dProgRgb, size_t(width) * height * 3));
            d->progW = width;
            d->progH = height;
        }
        SPT_CUDA_CHECK(cudaMemsetAsync(d->dProgAccum, 0, size_t(width) * height * sizeof(float4), d->stream));
    }
    return SPT_OK;
}

int spt_cuda_progressive_add(SptContext *c, const SptCamera *cam, const SptParams *prm, SptSegment seg)
{
    if (!c || !prm)
        return fail(SPT_ERR_INVALID, "null argument");
    int rc = check_camera(cam);
    if (rc)
        return rc;
    rc = check_segment(prm->width, prm->height, seg);
    if (rc)
        return rc;
    std::lock_guard<std::mutex> g(c->lock);
    if (!c->dProgAccum || prm->width != c->progW || prm->height != c->progH)
        return fail(SPT_ERR_INVALID, "spt_cuda_progressive_begin has not been called with this image size");
    if (c->all.size() > 1)
        return trace_multi(c, cam, prm, seg, [](SptContext *d) { return d->dProgAccum; });
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    return launch_trace(c, cam, prm, seg, c->dProgAccum);
}

int spt_cuda_progressive_image(SptContext *c, SptSegment seg, uint8_t *rgb8, float *accum4)
{
    if (!c)
        return fail(SPT_ERR_INVALID, "context is null");
    std::lock_guard<std::mutex> g(c->lock);
    if (!c->dProgAccum)
        return fail(SPT_ERR_INVALID, "spt_cuda_progressive_begin has not been called");
    uint32_t const W = c->progW, H = c->progH;
    int rc = check_segment(W, H, seg);
    if (rc)
        return rc;
    uint32_t const segW = seg.xEnd - seg.xBegin, segH = seg.yEnd - seg.yBegin;
    if (segW == 0 || segH == 0)
        return SPT_OK;
    if (c->all.size() > 1)
    {
        // the per-device sums go on accumulating afterwards: the totals are written to each device's scratch
        for (SptContext *d : c->all)
        {
            SPT_CUDA_CHECK(cudaSetDevice(d->device));
            rc = ensure_scratch(d, size_t(W) * H);
            if (rc)
                return rc;
        }
        return reduce_multi(c, W, H, seg, [](SptContext *d) { return d->dProgAccum; }, [](SptContext *d) { return d->dAccum; },
                            [](SptContext *d) { return d->dProgRgb; }, rgb8, accum4);
    }
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    if (rgb8)
    {
        rc = launch_resolve(c, c->dProgAccum, W, H, seg, c->dProgRgb);
        if (rc)
            return rc;
    }
    rc = copy_segment_to_host(c, W, H, seg, c->dProgRgb, rgb8, c->dProgAccum, accum4);
    if (rc)
        return rc;
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    return SPT_OK;
}

int spt_cuda_progressive_load(SptContext *c, const float *accum4)
{
    if (!c || !accum4)
        return fail(SPT_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->lock);
    if (!c->dProgAccum)
        return fail(SPT_ERR_INVALID, "spt_cuda_progressive_begin has not been called");
    // The checkpoint holds totals.  Every device continues from the rows of ITS bands of the full image (band b of 8 rows
    // belongs to device b % n, the ownership a full-image call with the band partition uses) and from zero elsewhere: a
    // resumed band-partitioned render continues every pixel's running sum on the device that owns it -- the bits of an
    // uninterrupted render -- and any other partition still sums to the same totals.
    uint32_t const n = uint32_t(c->all.size());
    for (uint32_t k = 0; k < n; ++k)
    {
        SptContext *d = c->all[k];
        SPT_CUDA_CHECK(cudaSetDevice(d->device));
        size_t const rowBytes = size_t(d->progW) * sizeof(float4);
        if (n == 1)
        {
            SPT_CUDA_CHECK(cudaMemcpyAsync(d->dProgAccum, accum4, rowBytes * d->progH, cudaMemcpyHostToDevice, d->stream));
        }
        else
        {
            SPT_CUDA_CHECK(cudaMemsetAsync(d->dProgAccum, 0, rowBytes * d->progH, d->stream));
            for (uint32_t y = k * SPT_BAND_ROWS; y < d->progH; y += n * SPT_BAND_ROWS)
            {
                uint32_t const rows = std::min<uint32_t>(SPT_BAND_ROWS, d->progH - y);
                SPT_CUDA_CHECK(cudaMemcpyAsync(d->dProgAccum + size_t(y) * d->progW, accum4 + size_t(y) * d->progW * 4, rowBytes * rows,
                                               cudaMemcpyHostToDevice, d->stream));
            }
        }
        SPT_CUDA_CHECK(cudaStreamSynchronize(d->stream));
    }
    return SPT_OK;
}

int spt_cuda_first_hit(SptContext *c, const SptCamera *cam, uint32_t width, uint32_t height, SptSegment seg, uint32_t flags,
                       uint32_t *idx, float *t, float *normal3)
{
    if (!c || !idx || !t || !normal3)
        return fail(SPT_ERR_INVALID, "null argument");
    int rc = check_camera(cam);
    if (rc)
        return rc;
    rc = check_segment(width, height, seg);
    if (rc)
        return rc;
    std::lock_guard<std::mutex> g(c->lock);
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    if (c->n == 0)
        return fail(SPT_ERR_NO_SCENE, "spt_cuda_set_scene has not been called");
    size_t const count = size_t(seg.xEnd - seg.xBegin) * (seg.yEnd - seg.yBegin);
    if (count == 0)
        return SPT_OK;
    rc = ensure_probe(c, count * 20);
    if (rc)
        return rc;
    spt::FirstHitParams p{};
    fill_camera(p.cam, cam);
    p.width = width;
    p.height = height;
    p.yBegin = seg.yBegin;
    p.yEnd = seg.yEnd;
    p.xBegin = seg.xBegin;
    p.xEnd = seg.xEnd;
    p.n = c->n;
    p.geom = c->dGeom;
    p.filter = c->filter;
    p.strict = (flags & SPT_FLAG_STRICT_LOOP) ? 1u : 0u;
    p.idx = static_cast<uint32_t *>(c->dProbe);
    p.t = reinterpret_cast<float *>(p.idx + count);
    p.normal3 = p.t + count;
    spt::first_hit_kernel<<<unsigned((count + 255) / 256), 256, 0, c->stream>>>(p);
    SPT_CUDA_CHECK(cudaGetLastError());
    SPT_CUDA_CHECK(cudaMemcpyAsync(idx, p.idx, count * 4, cudaMemcpyDeviceToHost, c->stream));
    SPT_CUDA_CHECK(cudaMemcpyAsync(t, p.t, count * 4, cudaMemcpyDeviceToHost, c->stream));
    SPT_CUDA_CHECK(cudaMemcpyAsync(normal3, p.normal3, count * 12, cudaMemcpyDeviceToHost, c->stream));
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    return SPT_OK;
}

int spt_cuda_closest_hit(SptContext *c, uint32_t count, const float *dir3, const float *org3, uint32_t flags, uint32_t *idx,
                         uint32_t *exact_evals)
{
    if (!c || !dir3 || !org3 || !idx)
        return fail(SPT_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->lock);
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    if (c->n == 0)
        return fail(SPT_ERR_NO_SCENE, "spt_cuda_set_scene has not been called");
    if (count == 0)
        return SPT_OK;
    int rc = ensure_probe(c, size_t(count) * 32);
    if (rc)
        return rc;
    float *dDir = static_cast<float *>(c->dProbe);
    float *dOrg = dDir + size_t(count) * 3;
    uint32_t *dIdx = reinterpret_cast<uint32_t *>(dOrg + size_t(count) * 3);
    uint32_t *dEvals = dIdx + count;
    uint32_t const mode = (flags & SPT_FLAG_STRICT_LOOP) ? 1u : ((flags & SPT_FLAG_CROSS_CHECK) ? 2u : 0u);
    SPT_CUDA_CHECK(cudaMemcpyAsync(dDir, dir3, size_t(count) * 12, cudaMemcpyHostToDevice, c->stream));
    SPT_CUDA_CHECK(cudaMemcpyAsync(dOrg, org3, size_t(count) * 12, cudaMemcpyHostToDevice, c->stream));
    bool useGrid = false;
    if (flags & SPT_FLAG_GRID) // the opt-in acceleration structure (spt_grid.cuh); with SPT_FLAG_CROSS_CHECK against the strict loop
    {
        rc = ensure_grid(c);
        if (rc)
            return rc;
        useGrid = c->gridUsable;
    }
    if (useGrid)
        spt::closest_hit_grid_kernel<<<(count + 255) / 256, 256, 0, c->stream>>>(c->dGeom, c->n, c->filter, c->grid, (flags & SPT_FLAG_CROSS_CHECK) ? 1u : 0u,
                                                                                   count, dDir, dOrg, dIdx, dEvals);
    else
        spt::closest_hit_kernel<<<(count + 255) / 256, 256, 0, c->stream>>>(c->dGeom, c->n, c->filter, mode, count, dDir, dOrg, dIdx, dEvals);
    SPT_CUDA_CHECK(cudaGetLastError());
    SPT_CUDA_CHECK(cudaMemcpyAsync(idx, dIdx, size_t(count) * 4, cudaMemcpyDeviceToHost, c->stream));
    if (exact_evals)
        SPT_CUDA_CHECK(cudaMemcpyAsync(exact_evals, dEvals, size_t(count) * 4, cudaMemcpyDeviceToHost, c->stream));
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    return SPT_OK;
}

int spt_cuda_closest_hit_static(SptContext *c, uint32_t count, const float *dir3, const float *org3, uint32_t flags, uint32_t *idx,
                                uint32_t *exact_evals)
{
    if (!c || !dir3 || !org3 || !idx)
        return fail(SPT_ERR_INVALID, "null argument");
    if (count > SPT_STATIC_SCRATCH_RAYS)
    {
        char msg[160];
        snprintf(msg, sizeof(msg), "spt_cuda_closest_hit_static: %u rays exceed the static scratch (SPT_STATIC_SCRATCH_RAYS = %u); split the call",
                 count, unsigned(SPT_STATIC_SCRATCH_RAYS));
        return fail(SPT_ERR_INVALID, msg);
    }
    std::lock_guard<std::mutex> g(c->lock);
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    if (!c->dStaticScratch)
        return fail(SPT_ERR_CUDA, "cudaGetSymbolAddress of the static scratch array failed at context creation", c->staticScratchError);
    if (c->n == 0)
        return fail(SPT_ERR_NO_SCENE, "spt_cuda_set_scene has not been called");
    if (count == 0)
        return SPT_OK;
    // the same carving as spt_cuda_closest_hit's probe buffer, inside the static array
    float *dDir = static_cast<float *>(c->dStaticScratch);
    float *dOrg = dDir + size_t(count) * 3;
    uint32_t *dIdx = reinterpret_cast<uint32_t *>(dOrg + size_t(count) * 3);
    uint32_t *dEvals = dIdx + count;
    uint32_t const mode = (flags & SPT_FLAG_STRICT_LOOP) ? 1u : ((flags & SPT_FLAG_CROSS_CHECK) ? 2u : 0u);
    SPT_CUDA_CHECK(cudaMemcpyAsync(dDir, dir3, size_t(count) * 12, cudaMemcpyHostToDevice, c->stream));
    SPT_CUDA_CHECK(cudaMemcpyAsync(dOrg, org3, size_t(count) * 12, cudaMemcpyHostToDevice, c->stream));
    SPT_CUDA_CHECK(cudaMemsetAsync(dIdx, 0xFF, size_t(count) * 8, c->stream)); // the array keeps the last call's answers: not these
    spt::closest_hit_kernel<<<(count + 255) / 256, 256, 0, c->stream>>>(c->dGeom, c->n, c->filter, mode, count, dDir, dOrg, dIdx,
                                                                          dEvals);
    SPT_CUDA_CHECK(cudaGetLastError());
    SPT_CUDA_CHECK(cudaMemcpyAsync(idx, dIdx, size_t(count) * 4, cudaMemcpyDeviceToHost, c->stream));
    if (exact_evals)
        SPT_CUDA_CHECK(cudaMemcpyAsync(exact_evals, dEvals, size_t(count) * 4, cudaMemcpyDeviceToHost, c->stream));
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    return SPT_OK;
}

int spt_cuda_get_counters(SptContext *c, SptCounters *out)
{
    if (!c || !out)
        return fail(SPT_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->lock);
    memset(out, 0, sizeof(*out));
    for (SptContext *d : c->all) // a multi-device context reports the sum over its devices
    {
        SPT_CUDA_CHECK(cudaSetDevice(d->device));
        unsigned long long h[8];
        SPT_CUDA_CHECK(cudaStreamSynchronize(d->stream));
        SPT_CUDA_CHECK(cudaMemcpy(h, d->dCounters, sizeof(h), cudaMemcpyDeviceToHost));
        out->segments += h[0];
        out->tests += h[0] * d->n;
        out->samples += h[1];
        out->dropped += h[2];
        out->exact += h[3];
        out->warp_iterations += h[4];
    }
    return SPT_OK;
}

int spt_cuda_reset_counters(SptContext *c)
{
    if (!c)
        return fail(SPT_ERR_INVALID, "context is null");
    std::lock_guard<std::mutex> g(c->lock);
    for (SptContext *d : c->all)
    {
        SPT_CUDA_CHECK(cudaSetDevice(d->device));
        SPT_CUDA_CHECK(cudaMemsetAsync(d->dCounters, 0, 8 * sizeof(unsigned long long), d->stream));
    }
    return SPT_OK;
}

int spt_cuda_get_timing(SptContext *c, SptTiming *out)
{
    if (!c || !out)
        return fail(SPT_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->lock);
    SptTiming t{};
    for (SptContext *d : c->all) // several devices work side by side: the slowest one's times, everybody's launches
    {
        SPT_CUDA_CHECK(cudaSetDevice(d->device));
        SPT_CUDA_CHECK(cudaStreamSynchronize(d->stream));
        float ms = 0.f;
        if (d->timing.launches && cudaEventElapsedTime(&ms, d->ev[0], d->ev[1]) == cudaSuccess)
            d->timing.trace_ms = ms;
        if (d->timedChunks)
        {
            SPT_CUDA_CHECK(cudaStreamSynchronize(d->auxStream));
            float pm = 0.f, sm = 0.f;
            for (uint32_t k = 0; k < d->timedChunks; ++k)
            {
                if (cudaEventElapsedTime(&ms, d->evKernel[4 * k + 0], d->evKernel[4 * k + 1]) == cudaSuccess)
                    pm += ms;
                if (cudaEventElapsedTime(&ms, d->evKernel[4 * k + 2], d->evKernel[4 * k + 3]) == cudaSuccess)
                    sm += ms;
            }
            d->timing.primary_ms = pm;
            d->timing.secondary_ms = sm;
        }
        if (cudaEventElapsedTime(&ms, d->ev[2], d->ev[3]) == cudaSuccess)
            d->timing.resolve_ms = ms;
        cudaGetLastError(); // events never recorded: not an error of the caller
        t.trace_ms = std::max(t.trace_ms, d->timing.trace_ms);
        t.resolve_ms = std::max(t.resolve_ms, d->timing.resolve_ms);
        t.launches += d->timing.launches;
        t.primary_ms = std::max(t.primary_ms, d->timing.primary_ms);
        t.secondary_ms = std::max(t.secondary_ms, d->timing.secondary_ms);
    }
    *out = t;
    return SPT_OK;
}

int spt_cuda_self_check_division(SptContext *c, uint64_t seed, uint64_t sets, uint64_t *tested, uint64_t *mismatches)
{
    if (!c || !tested || !mismatches)
        return fail(SPT_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->lock);
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    int rc = ensure_probe(c, 256);
    if (rc)
        return rc;
    unsigned long long *dOut = static_cast<unsigned long long *>(c->dProbe);
    SPT_CUDA_CHECK(cudaMemsetAsync(dOut, 0, 16, c->stream));
    uint32_t const threads = 256, blocks = uint32_t(c->prop.multiProcessorCount) * 8u;
    uint64_t const perThread = std::max<uint64_t>(1, (sets + uint64_t(threads) * blocks - 1) / (uint64_t(threads) * blocks));
    if (perThread > 0xFFFFFFFFull)
        return fail(SPT_ERR_INVALID, "too many operand sets");
    spt::division_selfcheck_kernel<<<blocks, threads, 0, c->stream>>>(seed, uint32_t(perThread), dOut);
    SPT_CUDA_CHECK(cudaGetLastError());
    unsigned long long h[2] = {0, 0};
    SPT_CUDA_CHECK(cudaMemcpyAsync(h, dOut, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    *tested = h[0];
    *mismatches = h[1];
    return SPT_OK;
}

int spt_cuda_measure_fp32_peak(SptContext *c, float *tflops_scalar, float *tflops_packed)
{
    if (!c)
        return fail(SPT_ERR_INVALID, "context is null");
    std::lock_guard<std::mutex> g(c->lock);
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    int rc = ensure_probe(c, 256);
    if (rc)
        return rc;
    float *dOut = static_cast<float *>(c->dProbe);
    int const blocks = c->prop.multiProcessorCount * 8, threads = 256;
    uint32_t const iters = 1u << 16;
    double const flopScalar = double(blocks) * threads * double(iters) * 8 * 2;
    float best[2] = {0.f, 0.f};
    for (int variant = 0; variant < 2; ++variant)
    {
        for (int rep = 0; rep < 4; ++rep)
        {
            SPT_CUDA_CHECK(cudaEventRecord(c->ev[0], c->stream));
            if (variant == 0)
                spt::fp32_peak_scalar<<<blocks, threads, 0, c->stream>>>(dOut, iters, 1.0000001f, 1e-9f);
            else
                spt::fp32_peak_packed<<<blocks, threads, 0, c->stream>>>(dOut, iters, 1.0000001f, 1e-9f);
            SPT_CUDA_CHECK(cudaGetLastError());
            SPT_CUDA_CHECK(cudaEventRecord(c->ev[1], c->stream));
            SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
            float ms = 0.f;
            SPT_CUDA_CHECK(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
            float const tf = float(flopScalar * (variant ? 2.0 : 1.0) / (double(ms) * 1e-3) / 1e12);
            if (rep > 0 && tf > best[variant])
                best[variant] = tf;
        }
    }
    if (tflops_scalar)
        *tflops_scalar = best[0];
    if (tflops_packed)
        *tflops_packed = best[1];
    return SPT_OK;
}

} // extern "C"
