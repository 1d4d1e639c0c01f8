// spt_cuda.cu -- C ABI (include/spt_cuda.h) over the sm_100a kernels.  No CPU fallback exists:
// every entry point returns an error when CUDA is unusable.
#include "../../include/spt_cuda.h"
#include "spt_kernels.cuh"
#include "spt_wavefront.cuh"
#include "spt_streamed.cuh"
#include "spt_split.cuh"
#include "spt_tensor_filter.cuh"

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

struct SptContext
{
    int device = 0;
    cudaDeviceProp prop{};
    cudaStream_t ownStream = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    std::mutex lock; // the reference calls its tracers from several host threads (Renderer.hpp:282-290)

    uint32_t n = 0;
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
    // split rendering: camera-form filter blob (rebuilt when the eye changes) and the first-hit staging buffer
    spt::FilterPair *dPairsCam = nullptr;
    std::vector<unsigned char> hostBlob; // the general blob, kept to derive the camera form
    std::vector<double> hostGeom;        // cx, cy, cz, r per SLOT in binary64
    float camEye[3] = {NAN, NAN, NAN};
    float camEpsT = 0.f;
    float4 *dFirstHits = nullptr;
    size_t firstHitsBytes = 0;
    // context-owned accumulator of the progressive entry points
    float4 *dProgAccum = nullptr;
    uint8_t *dProgRgb = nullptr;
    uint32_t progW = 0, progH = 0;
    void *dProbe = nullptr;
    size_t probeBytes = 0;
    size_t maxSmemOptin = 0;
#if SPT_TENSOR_FILTER
    float *dTensorB = nullptr; // sphere matrix of the tensor-core filter (spt_tensor_host.hpp)
    float *dTensorBCam = nullptr; // its camera form, rebuilt with the camera blob when the eye changes
    uint32_t tensorRows = 0;
#endif
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

// Camera form of the blob (spt_filter.cuh, filter_pair_cam): per slot nkc = -(|C - E|^2 - r^2) + S_i + S_ray(E) - 1e-3 in
// binary64, rounded up; S_ray and epsT exactly as make_filter_ray derives them for an origin at E.
int ensure_camera_blob(SptContext *c, const float eye[3])
{
    if (c->dPairsCam && eye[0] == c->camEye[0] && eye[1] == c->camEye[1] && eye[2] == c->camEye[2])
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
    for (uint32_t k = 0; k < nSlots; ++k)
    {
        double const cx = c->hostGeom[4 * k], cy = c->hostGeom[4 * k + 1], cz = c->hostGeom[4 * k + 2], r = c->hostGeom[4 * k + 3];
        if (!(r >= 0.0)) // padding slot: stays never-pass
            continue;
        double const a2 = cx * cx + cy * cy + cz * cz;
        double const dx = cx - ex, dy = cy - ey, dz = cz - ez;
        double const slack = spt::kSlackU * (a2 + r * r);
        pairs[k / 2].nk[k & 1] = round_up_to_float(-(dx * dx + dy * dy + dz * dz - r * r) + slack + sRay - double(1e-3f));
    }
    if (!c->dPairsCam)
        SPT_CUDA_CHECK(cudaMalloc(&c->dPairsCam, blob.size() + 16));
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream)); // an earlier launch may still read the previous camera's records
    if (!blob.empty())
        SPT_CUDA_CHECK(cudaMemcpy(c->dPairsCam, blob.data(), blob.size(), cudaMemcpyHostToDevice));
#if SPT_TENSOR_FILTER
    if (nSlots > 0)
    {
        std::vector<float> B;
        double const e[3] = {ex, ey, ez};
        spt::build_tensor_matrix_cam(c->hostGeom, nSlots, e, B);
        if (!c->dTensorBCam)
            SPT_CUDA_CHECK(cudaMalloc(&c->dTensorBCam, B.size() * sizeof(float)));
        SPT_CUDA_CHECK(cudaMemcpy(c->dTensorBCam, B.data(), B.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
#endif
    c->camEye[0] = eye[0];
    c->camEye[1] = eye[1];
    c->camEye[2] = eye[2];
    c->camEpsT = round_up_to_float(epsT);
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
        if (geomBytes + pairBytes <= wfBudget)
            SPT_LAUNCH_WF(2, geomBytes + pairBytes);
        else if (pairBytes <= wfBudget)
            SPT_LAUNCH_WF(1, pairBytes);
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
            int rc = ensure_camera_blob(c, cam->eye);
            if (rc)
                return rc;
            p.filterCam = c->filter;
            p.filterCam.pairs = c->dPairsCam;
            p.filterCam.slotIdx = reinterpret_cast<const uint32_t *>(c->dPairsCam + c->filter.nSlots / 2);
            p.filterCam.epsTCam = c->camEpsT;
            size_t const pixels = size_t(segW) * std::min<size_t>(segH, size_t(myBands) * SPT_BAND_ROWS); // rows the launch covers
            // Staging budget.  A lane owns its pixel for a whole chunk and the warp waits for its slowest lane, so longer
            // chunks waste fewer lane-iterations (config 3 on B200: 8 / 16 / 32 / 64 GiB = 5 815 / 5 893 / 5 963 / 6 006
            // Msamples/s, profiles/r1_ab_chunk_and_cta_shape.txt): up to 64 GiB (512 spp of a 3840x2160 render), but never
            // more than 45 % of what the device has free.
            size_t cap = size_t(64) << 30;
            {
                size_t freeB = 0, totalB = 0;
                if (cudaMemGetInfo(&freeB, &totalB) == cudaSuccess)
                    cap = std::min(cap, (freeB + c->firstHitsBytes) / 100 * 45);
                else
                    (void)cudaGetLastError();
            }
            if (const char *knob = std::getenv("SPT_SPLIT_CAP_MIB")) // experiments: staging budget in MiB
                cap = std::max<size_t>(size_t(1) << 20, size_t(std::atoll(knob)) << 20);
            uint32_t const total = prm->sample_end - prm->sample_begin;
            uint32_t chunk = uint32_t(std::max<size_t>(1, std::min<size_t>(total, cap / (pixels * sizeof(float4)))));
            if (const char *knob = std::getenv("SPT_SPLIT_CHUNK")) // test knob: force several chunks on small renders
                chunk = std::max<uint32_t>(1u, std::min<uint32_t>(chunk, uint32_t(std::atoi(knob))));
            size_t const pad = 256; // the consumer prefetches four records past the one it reads
            if (pixels * chunk * sizeof(float4) + pad > c->firstHitsBytes)
            {
                SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
                cudaFree(c->dFirstHits);
                c->dFirstHits = nullptr;
                c->firstHitsBytes = 0;
                // a device short of memory gets smaller chunks (more launches), not an error
                for (;;)
                {
                    size_t const need = pixels * chunk * sizeof(float4) + pad;
                    cudaError_t const ce = cudaMalloc(&c->dFirstHits, need);
                    if (ce == cudaSuccess)
                    {
                        c->firstHitsBytes = need;
                        break;
                    }
                    (void)cudaGetLastError();
                    c->dFirstHits = nullptr;
                    if (chunk == 1)
                        return fail(SPT_ERR_CUDA, "cudaMalloc of the first-hit staging buffer failed", ce);
                    chunk = (chunk + 1) / 2;
                }
            }
            p.firstHits = c->dFirstHits;
            void (*primary)(const spt::TraceParams) = streamed ? spt::primary_hits_streamed : spt::primary_hits_kernel<2, spt::SplitShape>;
            void (*secondary)(const spt::TraceParams) = streamed ? spt::trace_secondary_streamed : spt::trace_secondary<2, spt::SplitShape>;
            // the resident kernels run smaller CTAs than the rest (spt_split.cuh, SplitShape)
            dim3 const sGrid = streamed ? grid : dim3((segW + spt::SplitShape::kTileW - 1) / spt::SplitShape::kTileW, myBands);
            dim3 const sBlock(streamed ? spt::kBlockThreads : spt::SplitShape::kThreads);
            SPT_CUDA_CHECK(cudaFuncSetAttribute(primary, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            SPT_CUDA_CHECK(cudaFuncSetAttribute(secondary, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            size_t smem2 = smem, smem1 = smem;
            dim3 sGrid2 = sGrid, sBlock2 = sBlock, sGrid1 = sGrid, sBlock1 = sBlock;
#if SPT_TENSOR_FILTER
            // experimental build: the resident bounce kernel takes its discriminants from the tensor core unless SPT_TENSOR=0
            {
                const char *knob = std::getenv("SPT_TENSOR");
                if (!streamed && c->dTensorB && !(knob && knob[0] == '0'))
                {
                    p.tensorB = c->dTensorB;
                    p.tensorRows = c->tensorRows;
                    secondary = spt::trace_secondary_tc;
                    sGrid2 = dim3((segW + spt::TensorShape::kTileW - 1) / spt::TensorShape::kTileW, myBands);
                    sBlock2 = dim3(spt::TensorShape::kThreads);
                    smem2 = ((geomBytes + pairBytes + 127) & ~size_t(127)) + spt::tensor_b_bytes(c->tensorRows) +
                            size_t(spt::kTensorAcc) * spt::kTensorABytes;
                    SPT_CUDA_CHECK(cudaFuncSetAttribute(secondary, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem2)));
                }
                // ... and the camera kernel unless SPT_TENSOR_PRIMARY=0
                knob = std::getenv("SPT_TENSOR_PRIMARY");
                if (!streamed && c->dTensorBCam && !(knob && knob[0] == '0'))
                {
                    p.tensorBCam = c->dTensorBCam;
                    p.tensorRows = c->tensorRows;
                    primary = spt::primary_hits_tc;
                    sGrid1 = dim3((segW + spt::TensorShape::kTileW - 1) / spt::TensorShape::kTileW, myBands);
                    sBlock1 = dim3(spt::TensorShape::kThreads);
                    smem1 = ((geomBytes + pairBytes + 127) & ~size_t(127)) + spt::tensor_b_bytes_cam(c->tensorRows) +
                            size_t(spt::kTensorAcc) * spt::kTensorABytesCam;
                    SPT_CUDA_CHECK(cudaFuncSetAttribute(primary, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem1)));
                }
            }
#endif
            for (uint32_t s0 = prm->sample_begin; s0 < prm->sample_end; s0 += chunk)
            {
                p.sampleBegin = s0;
                p.sampleEnd = std::min(prm->sample_end, s0 + chunk);
                primary<<<sGrid1, sBlock1, smem1, c->stream>>>(p);
                secondary<<<sGrid2, sBlock2, smem2, c->stream>>>(p);
                c->timing.launches += 2;
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
    for (int i = 0; i < 4 && ce == cudaSuccess; ++i)
        ce = cudaEventCreate(&c->ev[i]);
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
    *out = c;
    return SPT_OK;
}

void spt_cuda_destroy(SptContext *c)
{
    if (!c)
        return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    cudaFree(c->dGeom);
    cudaFree(c->dShade);
    cudaFree(c->dMat);
    cudaFree(c->dPairs);
    cudaFree(c->dBigIdx);
    cudaFree(c->dCounters);
    cudaFree(c->dAccum);
    cudaFree(c->dRgb);
    cudaFree(c->dProbe);
    cudaFree(c->dProgAccum);
    cudaFree(c->dProgRgb);
    cudaFree(c->dPairsCam);
    cudaFree(c->dFirstHits);
#if SPT_TENSOR_FILTER
    cudaFree(c->dTensorB);
    cudaFree(c->dTensorBCam);
#endif
    for (auto &e : c->ev)
        if (e)
            cudaEventDestroy(e);
    if (c->ownStream)
        cudaStreamDestroy(c->ownStream);
    delete c;
}

int spt_cuda_set_stream(SptContext *c, void *cuda_stream, int use_own)
{
    if (!c)
        return fail(SPT_ERR_INVALID, "context is null");
    std::lock_guard<std::mutex> g(c->lock);
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
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    std::vector<float4> geom(s->n), shade(s->n);
    for (uint32_t i = 0; i < s->n; ++i)
    {
        if (s->centers4[4 * i + 3] != 0.f)
            return fail(SPT_ERR_INVALID, "sphere centre w must be 0 (Vec4{x,y,z} initialisers, SceneGenerators.hpp)");
        geom[i] = make_float4(s->centers4[4 * i], s->centers4[4 * i + 1], s->centers4[4 * i + 2], s->radii[i]);
        shade[i] = make_float4(s->colors4[4 * i], s->colors4[4 * i + 1], s->colors4[4 * i + 2], s->fuzz[i]);
    }
    std::vector<unsigned char> blob;
    std::vector<uint32_t> big;
    uint32_t nSlots = 0;
    float amax = 0.f;
    build_filter(s, blob, nSlots, big, amax);
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    cudaFree(c->dGeom);
    cudaFree(c->dShade);
    cudaFree(c->dMat);
    cudaFree(c->dPairs);
    cudaFree(c->dBigIdx);
    c->dGeom = c->dShade = nullptr;
    c->dMat = nullptr;
    c->dPairs = nullptr;
    c->dBigIdx = nullptr;
    c->n = 0;
    SPT_CUDA_CHECK(cudaMalloc(&c->dPairs, blob.size() + 16));
    SPT_CUDA_CHECK(cudaMalloc(&c->dBigIdx, (big.size() + 1) * sizeof(uint32_t)));
    if (!blob.empty())
        SPT_CUDA_CHECK(cudaMemcpy(c->dPairs, blob.data(), blob.size(), cudaMemcpyHostToDevice));
    if (!big.empty())
        SPT_CUDA_CHECK(cudaMemcpy(c->dBigIdx, big.data(), big.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    // host copies from which the camera form of the records is derived per eye (ensure_camera_blob)
    c->hostBlob = blob;
    c->hostGeom.assign(size_t(nSlots) * 4, -1.0);
    {
        const uint32_t *slotIdx = reinterpret_cast<const uint32_t *>(blob.data() + size_t(nSlots) / 2 * sizeof(spt::FilterPair));
        for (uint32_t k = 0; k < nSlots; ++k)
        {
            uint32_t const i = slotIdx[k];
            if (i >= s->n)
                continue; // padding: radius stays -1
            c->hostGeom[4 * k + 0] = s->centers4[4 * i];
            c->hostGeom[4 * k + 1] = s->centers4[4 * i + 1];
            c->hostGeom[4 * k + 2] = s->centers4[4 * i + 2];
            c->hostGeom[4 * k + 3] = std::fabs(double(s->radii[i]));
        }
    }
#if SPT_TENSOR_FILTER
    cudaFree(c->dTensorB);
    cudaFree(c->dTensorBCam); // sized by the scene; ensure_camera_blob rebuilds it (camEye is reset below)
    c->dTensorB = c->dTensorBCam = nullptr;
    c->tensorRows = spt::tensor_rows(nSlots);
    if (nSlots > 0)
    {
        std::vector<float> B;
        spt::build_tensor_matrix(c->hostGeom, nSlots, B);
        SPT_CUDA_CHECK(cudaMalloc(&c->dTensorB, B.size() * sizeof(float)));
        SPT_CUDA_CHECK(cudaMemcpy(c->dTensorB, B.data(), B.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
#endif
    cudaFree(c->dPairsCam);
    c->dPairsCam = nullptr;
    c->camEye[0] = c->camEye[1] = c->camEye[2] = NAN;
    c->filter.pairs = c->dPairs;
    c->filter.nSlots = nSlots;
    c->filter.slotIdx = reinterpret_cast<const uint32_t *>(c->dPairs + nSlots / 2);
    c->filter.bigIdx = c->dBigIdx;
    c->filter.nBig = uint32_t(big.size());
    c->filter.amax = amax;
    SPT_CUDA_CHECK(cudaMalloc(&c->dGeom, s->n * sizeof(float4)));
    SPT_CUDA_CHECK(cudaMalloc(&c->dShade, s->n * sizeof(float4)));
    SPT_CUDA_CHECK(cudaMalloc(&c->dMat, s->n));
    SPT_CUDA_CHECK(cudaMemcpy(c->dGeom, geom.data(), s->n * sizeof(float4), cudaMemcpyHostToDevice));
    SPT_CUDA_CHECK(cudaMemcpy(c->dShade, shade.data(), s->n * sizeof(float4), cudaMemcpyHostToDevice));
    SPT_CUDA_CHECK(cudaMemcpy(c->dMat, s->materials, s->n, cudaMemcpyHostToDevice));
    c->n = s->n;
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
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
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
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    uint32_t const W = prm->width, H = prm->height;
    uint32_t const segW = seg.xEnd - seg.xBegin, segH = seg.yEnd - seg.yBegin;
    if (segW == 0 || segH == 0)
        return SPT_OK;
    rc = ensure_scratch(c, size_t(W) * H);
    if (rc)
        return rc;
    // zero the accumulator rows of the segment
    SPT_CUDA_CHECK(cudaMemset2DAsync(c->dAccum + size_t(seg.yBegin) * W + seg.xBegin, size_t(W) * sizeof(float4), 0,
                                     size_t(segW) * sizeof(float4), segH, c->stream));
    rc = launch_trace(c, cam, prm, seg, c->dAccum);
    if (rc)
        return rc;
    uint32_t const traceLaunches = c->timing.launches;
    rc = launch_resolve(c, c->dAccum, W, H, seg, c->dRgb);
    if (rc)
        return rc;
    c->timing.launches += traceLaunches;
    // pixel (x,y) lives in byte row H-1-y: copy the segment's rows back into the caller's g_data
    uint32_t const rowTop = H - seg.yEnd; // flipped
    SPT_CUDA_CHECK(cudaMemcpy2DAsync(rgb8 + (size_t(rowTop) * W + seg.xBegin) * 3, size_t(W) * 3,
                                     c->dRgb + (size_t(rowTop) * W + seg.xBegin) * 3, size_t(W) * 3, size_t(segW) * 3, segH,
                                     cudaMemcpyDeviceToHost, c->stream));
    if (accum4)
        SPT_CUDA_CHECK(cudaMemcpy2DAsync(accum4 + (size_t(seg.yBegin) * W + seg.xBegin) * 4, size_t(W) * sizeof(float4),
                                         c->dAccum + size_t(seg.yBegin) * W + seg.xBegin, size_t(W) * sizeof(float4),
                                         size_t(segW) * sizeof(float4), segH, cudaMemcpyDeviceToHost, c->stream));
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
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    if (width != c->progW || height != c->progH || !c->dProgAccum)
    {
        cudaFree(c->dProgAccum);
        cudaFree(c->dProgRgb);
        c->dProgAccum = nullptr;
        c->dProgRgb = nullptr;
        c->progW = c->progH = 0;
        SPT_CUDA_CHECK(cudaMalloc(&c->dProgAccum, size_t(width) * height * sizeof(float4)));
        SPT_CUDA_CHECK(cudaMalloc(&c->dProgRgb, size_t(width) * height * 3));
        c->progW = width;
        c->progH = height;
    }
    SPT_CUDA_CHECK(cudaMemsetAsync(c->dProgAccum, 0, size_t(width) * height * sizeof(float4), c->stream));
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
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    if (!c->dProgAccum || prm->width != c->progW || prm->height != c->progH)
        return fail(SPT_ERR_INVALID, "spt_cuda_progressive_begin has not been called with this image size");
    return launch_trace(c, cam, prm, seg, c->dProgAccum);
}

int spt_cuda_progressive_image(SptContext *c, SptSegment seg, uint8_t *rgb8, float *accum4)
{
    if (!c)
        return fail(SPT_ERR_INVALID, "context is null");
    std::lock_guard<std::mutex> g(c->lock);
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    if (!c->dProgAccum)
        return fail(SPT_ERR_INVALID, "spt_cuda_progressive_begin has not been called");
    uint32_t const W = c->progW, H = c->progH;
    int rc = check_segment(W, H, seg);
    if (rc)
        return rc;
    uint32_t const segW = seg.xEnd - seg.xBegin, segH = seg.yEnd - seg.yBegin;
    if (segW == 0 || segH == 0)
        return SPT_OK;
    if (rgb8)
    {
        rc = launch_resolve(c, c->dProgAccum, W, H, seg, c->dProgRgb);
        if (rc)
            return rc;
        uint32_t const rowTop = H - seg.yEnd; // pixel (x,y) lives in byte row H-1-y
        SPT_CUDA_CHECK(cudaMemcpy2DAsync(rgb8 + (size_t(rowTop) * W + seg.xBegin) * 3, size_t(W) * 3,
                                         c->dProgRgb + (size_t(rowTop) * W + seg.xBegin) * 3, size_t(W) * 3, size_t(segW) * 3, segH,
                                         cudaMemcpyDeviceToHost, c->stream));
    }
    if (accum4)
        SPT_CUDA_CHECK(cudaMemcpy2DAsync(accum4 + (size_t(seg.yBegin) * W + seg.xBegin) * 4, size_t(W) * sizeof(float4),
                                         c->dProgAccum + size_t(seg.yBegin) * W + seg.xBegin, size_t(W) * sizeof(float4),
                                         size_t(segW) * sizeof(float4), segH, cudaMemcpyDeviceToHost, c->stream));
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    return SPT_OK;
}

int spt_cuda_progressive_load(SptContext *c, const float *accum4)
{
    if (!c || !accum4)
        return fail(SPT_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->lock);
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    if (!c->dProgAccum)
        return fail(SPT_ERR_INVALID, "spt_cuda_progressive_begin has not been called");
    SPT_CUDA_CHECK(cudaMemcpyAsync(c->dProgAccum, accum4, size_t(c->progW) * c->progH * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
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
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    unsigned long long h[8];
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    SPT_CUDA_CHECK(cudaMemcpy(h, c->dCounters, sizeof(h), cudaMemcpyDeviceToHost));
    out->segments = h[0];
    out->tests = h[0] * c->n;
    out->samples = h[1];
    out->dropped = h[2];
    out->exact = h[3];
    return SPT_OK;
}

int spt_cuda_reset_counters(SptContext *c)
{
    if (!c)
        return fail(SPT_ERR_INVALID, "context is null");
    std::lock_guard<std::mutex> g(c->lock);
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    SPT_CUDA_CHECK(cudaMemsetAsync(c->dCounters, 0, 8 * sizeof(unsigned long long), c->stream));
    return SPT_OK;
}

int spt_cuda_get_timing(SptContext *c, SptTiming *out)
{
    if (!c || !out)
        return fail(SPT_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->lock);
    SPT_CUDA_CHECK(cudaSetDevice(c->device));
    SPT_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]) == cudaSuccess)
        c->timing.trace_ms = ms;
    if (cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]) == cudaSuccess)
        c->timing.resolve_ms = ms;
    cudaGetLastError(); // events never recorded: not an error of the caller
    *out = c->timing;
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
