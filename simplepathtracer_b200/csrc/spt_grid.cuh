// spt_grid.cuh -- SURVEY 8(f)-4: an acceleration structure for long sphere lists (the 16k-sphere case), OPT-IN (SPT_FLAG_GRID).
//
// The north star's path is the reference's LINEAR closest hit (Collision.hpp:87-109) and everything the roofline counts is
// defined on it; this file is the "what comes next" row, kept apart from it: a uniform grid over the filterable spheres that
// a lane walks cell by cell along its ray (3-D DDA), running the reference's exact sequence (exact_sphere) on the spheres
// listed in each cell.  The result is FindClosestIntersectionSphere's, bit for bit -- the exact sequence alone decides index
// and distance, ties resolve on (distance^2, index) = "lowest index wins" (Collision.hpp:102), spheres containing the origin
// are ignored by the sequence itself -- because the structure only ever SKIPS spheres that provably cannot pass:
//
//   * what "can pass" means in float.  The sequence passes when its float discriminant r^2 - (|L|^2 - tC^2) > 1e-3.  Its
//     distance from the real discriminant is at most E = u (33.4 (A^2 + B^2) + 2 r^2) (u = 2^-24, A = |C|, B = |O|; DESIGN.md,
//     "pre-filter bound"), so the real LINE passes within sqrt(r^2 + E) of the centre, and the float contact point lies within
//     that distance (+ rounding of the point) of the centre as well.  The host lists sphere i in every cell its box of half
//     width R_i = sqrt(r_i^2 + E_i(Bmax)) + m overlaps (42 in place of 33.4; m below), for origins with |O| <= Bmax.
//   * origins beyond Bmax (contact points far out on the ground sphere) do not use the cells: E grows with B^2, so their test
//     is against the grid's box fattened by sqrt(rmax^2 + E(B)) + m; a ray that misses it cannot pass any listed sphere, one
//     that does not goes through all of them (the slot map of the pre-filter), exactly as the strict loop would.
//   * the walk in float.  A computed boundary crossing is off by at most ~2^-23 (|coordinate| + |origin|) in the crossing
//     COORDINATE whatever the direction (the error in t is that over |d_a|: large for rays nearly parallel to a plane, and
//     harmless, because such a ray stays that close to the plane).  So the cells visited differ from the cells the real line
//     crosses only within that distance of a cell boundary, and m >= 2^-16 (Amax + Bmax) >> it: a sphere the ray can touch there
//     is listed on both sides.
//   * when to stop.  After the cell that the ray leaves at tNext, every sphere not yet tested has its contact point beyond
//     tNext; its float distance^2 is at least ((tNext - a)(1 - 2^-16))^2 with a = 2^-20 (Bmax + extent) for the rounding of the
//     contact point.  If the best hit so far (listed spheres and the big ones, e.g. the ground) is closer than that, nothing
//     later can win or tie.
//
// Spheres the pre-filter keeps out (the ground: r > 64 or |C| > 4096) stay out of the grid too and are evaluated first.
#pragma once

#include "spt_filter.cuh"

namespace spt
{

struct GridScene
{
    const uint32_t *cellStart; // nCells + 1 offsets into refs; cell (ix, iy, iz) = (iz * dim[1] + iy) * dim[0] + ix
    const uint32_t *refs;      // sphere indices
    float lo[3], hi[3];        // the grid's box
    float size[3], inv[3];     // cell size per axis and its reciprocal
    uint32_t dim[3];
    float bmaxSq;              // origins with |O|^2 above this do not use the cells
    float fatU;                // 42 u: far origins, fattening = sqrt(rmaxSq + fatU (amaxSq + |O|^2)) + margin
    float rmaxSq, amaxSq, margin;
    float stopAbs;             // a of the stopping rule
};

// every filterable sphere through the exact sequence: what a far origin whose ray reaches the fattened box falls back to
template <typename GeomPtr>
__device__ __forceinline__ void exact_all_slots(const FilterScene &fs, GeomPtr geom, uint32_t n, V3 dir, V3 org, float od, uint32_t &minIndex,
                                                float &minDistanceSq, uint32_t &exactEvals)
{
    for (uint32_t k = 0; k < fs.nSlots; ++k)
    {
        uint32_t const i = fs.slotIdx[k];
        if (i < n)
        {
            exact_sphere<false>(geom[i], i, dir, org, od, minIndex, minDistanceSq);
            ++exactEvals;
        }
    }
}

// ray (t >= 0) against the box [lo - pad, hi + pad]: parameter interval, empty when tEnter > tExit (NaN: empty)
__device__ __forceinline__ bool clip_to_box(const GridScene &g, float pad, V3 dir, V3 org, float &tEnter, float &tExit)
{
    float const o[3] = {org.x, org.y, org.z}, d[3] = {dir.x, dir.y, dir.z};
    tEnter = 0.f;
    tExit = FLT_MAX;
#pragma unroll
    for (int a = 0; a < 3; ++a)
    {
        float const l = g.lo[a] - pad, h = g.hi[a] + pad;
        if (fabsf(d[a]) < 1e-20f)
        {
            if (!(o[a] >= l && o[a] <= h))
                return false;
        }
        else
        {
            float const r = 1.f / d[a];
            float const t0 = (l - o[a]) * r, t1 = (h - o[a]) * r;
            tEnter = fmaxf(tEnter, fminf(t0, t1));
            tExit = fminf(tExit, fmaxf(t0, t1));
        }
    }
    return tEnter <= tExit;
}

// FindClosestIntersectionSphere (Collision.hpp:87-109) through the grid.  Per lane, no warp collectives inside: lanes walk
// different numbers of cells.  `fs` supplies the big spheres and, for the fallback, the list of filterable ones.
template <typename GeomPtr>
__device__ __forceinline__ uint32_t closest_hit_grid(const GridScene &g, const FilterScene &fs, GeomPtr geom, uint32_t n, V3 dir, V3 org,
                                                     uint32_t &exactEvals, bool live)
{
    uint32_t minIndex = n;
    float minDistanceSq = FLT_MAX;
    if (!live)
        return minIndex;
    float const od = vdot(org, dir);
    exact_big_spheres(fs, geom, dir, org, od, minIndex, minDistanceSq, exactEvals, true);
    float const oo = __fmaf_ru(org.z, org.z, __fmaf_ru(org.y, org.y, __fmul_ru(org.x, org.x)));
    float tEnter, tExit;
    if (!(oo <= g.bmaxSq))
    {
        // far origin (or NaN): no cells; either the ray stays clear of everything listed, or every listed sphere is tried
        float const fat = __fadd_ru(__fsqrt_ru(__fmaf_ru(g.fatU, __fadd_ru(g.amaxSq, oo), g.rmaxSq)), g.margin);
        if (!(oo < FLT_MAX) || clip_to_box(g, fat, dir, org, tEnter, tExit))
            exact_all_slots(fs, geom, n, dir, org, od, minIndex, minDistanceSq, exactEvals);
        return minIndex;
    }
    if (!clip_to_box(g, 0.f, dir, org, tEnter, tExit))
        return minIndex;

    // per axis: cell index, direction of travel, reciprocal of the direction (0 for an axis the ray does not move along: its
    // crossing time is never the smallest).  Crossing times are recomputed from the boundary's coordinate at every step (no
    // running sums), scalars throughout (an array indexed by the stepping axis would live in local memory).
    struct Axis
    {
        float o, r, lo, size, tMax;
        int cell, step, dim;
        __device__ __forceinline__ void init(float org, float dir, float tEnter, float gLo, float gInv, float gSize, uint32_t gDim)
        {
            o = org;
            lo = gLo;
            size = gSize;
            dim = int(gDim);
            cell = max(0, min(dim - 1, int(floorf((fmaf(dir, tEnter, org) - gLo) * gInv))));
            bool const moves = fabsf(dir) >= 1e-20f;
            r = moves ? 1.f / dir : 0.f;
            step = !moves ? 0 : (dir > 0.f ? 1 : -1);
            crossing();
        }
        __device__ __forceinline__ void crossing() { tMax = step == 0 ? FLT_MAX : (fmaf(float(cell + (step > 0 ? 1 : 0)), size, lo) - o) * r; }
        __device__ __forceinline__ bool advance() // false: the ray leaves the grid
        {
            cell += step;
            if (cell < 0 || cell >= dim)
                return false;
            crossing();
            return true;
        }
    };
    Axis ax, ay, az;
    ax.init(org.x, dir.x, tEnter, g.lo[0], g.inv[0], g.size[0], g.dim[0]);
    ay.init(org.y, dir.y, tEnter, g.lo[1], g.inv[1], g.size[1], g.dim[1]);
    az.init(org.z, dir.z, tEnter, g.lo[2], g.inv[2], g.size[2], g.dim[2]);
    uint32_t const maxSteps = g.dim[0] + g.dim[1] + g.dim[2] + 3u;
    for (uint32_t s = 0; s < maxSteps; ++s)
    {
        uint32_t const c = (uint32_t(az.cell) * g.dim[1] + uint32_t(ay.cell)) * g.dim[0] + uint32_t(ax.cell);
        uint32_t const k1 = __ldg(g.cellStart + c + 1);
        for (uint32_t k = __ldg(g.cellStart + c); k < k1; ++k)
        {
            uint32_t const i = __ldg(g.refs + k);
            exact_sphere<false>(geom[i], i, dir, org, od, minIndex, minDistanceSq);
            ++exactEvals;
        }
        // the ray leaves the cell through the boundary that comes first, at tNext
        float const tNext = fminf(ax.tMax, fminf(ay.tMax, az.tMax));
        float const safe = fmaxf(0.f, (tNext - g.stopAbs) * 0.99998474f); // 1 - 2^-16
        if (minDistanceSq < safe * safe || !(tNext <= tExit))
            break;
        bool inGrid;
        if (ax.tMax <= ay.tMax && ax.tMax <= az.tMax)
            inGrid = ax.advance();
        else if (ay.tMax <= az.tMax)
            inGrid = ay.advance();
        else
            inGrid = az.advance();
        if (!inGrid)
            break;
    }
    return minIndex;
}

// parity probe (spt_cuda_closest_hit with SPT_FLAG_GRID): caller-supplied rays through the grid; crossCheck also runs the strict loop
// and reports 0xFFFFFFFF where the two disagree (must never happen)
__global__ void closest_hit_grid_kernel(const float4 *geom, uint32_t n, FilterScene filter, GridScene grid, uint32_t crossCheck, uint32_t count,
                                        const float *dir3, const float *org3, uint32_t *idx, uint32_t *evals)
{
    uint32_t const k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count)
        return;
    V3 const d = v3(dir3[3 * k], dir3[3 * k + 1], dir3[3 * k + 2]);
    V3 const o = v3(org3[3 * k], org3[3 * k + 1], org3[3 * k + 2]);
    uint32_t exactEvals = 0;
    uint32_t r = closest_hit_grid(grid, filter, geom, n, d, o, exactEvals, true);
    if (crossCheck && r != closest_hit_strict(geom, n, d, o))
        r = 0xFFFFFFFFu;
    idx[k] = r;
    if (evals)
        evals[k] = exactEvals;
}

} // namespace spt
