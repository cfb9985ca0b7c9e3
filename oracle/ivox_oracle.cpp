// ============================================================================
// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// CPU restatement of the iVox branch of the hot path: the structure the reference
// builds when USE_ikdtree is off (CMakeLists.txt:15), faster_lio::IVox<3, IVoxNodeType::DEFAULT, PointType>
// (include/laser_mapping.h:65).  All citations are relative to /root/reference/.
//
// PARITY: PINNED ON THE CONTAINER LOGIC, not on three Eigen calls.  The reference holds no test or golden vector for
// this branch, and ivox3d.h needs Eigen, glog and pcl, which are not in this image.  Its headers do compile UNMODIFIED
// against the small stand-ins of oracle/shim (Eigen column vectors, pcl point types, glog): oracle/ivox_ref_wrap.cpp ->
// oracle/_ref/libivox_ref.so, and tests/test_ivox_cpu.py holds this file to that library for all four stencils — voxel
// count, found / not found, neighbour count, neighbour set, nearest element, LRU eviction at capacity_.  What stays a
// restatement (the stand-in and this file share my reading of Eigen) is the arithmetic of squaredNorm, array().round()
// and cast<int>():
//   * Pos2Grid (ivox3d.h:278-281): (pt * inv_resolution_).array().round().cast<int>(), inv_resolution_ a float
//     holding 1.0 / resolution_ (ivox3d.h:56,63); std::round semantics (half away from zero).
//   * distance2 (ivox3d_node.hpp:13-16): (pt1 - pt2).squaredNorm() on Vector3f; Eigen's unrolled reduction of a
//     3-vector is x*x + (y*y + z*z).
//   * GetClosestPoint (ivox3d.h:132-204) leaves its result in the order two std::nth_element calls produce (only
//     element 0 is the minimum).  Here the list is ascending under (distance, insertion id): the same SET, a
//     canonical order.  Equal distances at rank K are resolved by insertion id instead of nth_element's choice;
//     the tie counter lets tests assert there are none in the data.
//   * LRU eviction at Options::capacity_ voxels (ivox3d.h:262-265, default 1 000 000): restated (see add_points),
//     so tests can show the product's documented deviation (no eviction) does not matter below capacity.
// ============================================================================
#include "flivo_oracle.h"

#include <math.h>
#include <algorithm>
#include <array>
#include <list>
#include <unordered_map>
#include <vector>

namespace {

struct KeyHash3 {   // any hash works: iteration order of the container is never observed
    size_t operator()(const std::array<int32_t, 3>& k) const {
        uint64_t h = (uint64_t)(uint32_t)k[0] * 0x9E3779B97F4A7C15ull;
        h ^= ((uint64_t)(uint32_t)k[1] + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full + (h << 6) + (h >> 2);
        h ^= ((uint64_t)(uint32_t)k[2] + 0x165667B1ull) * 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
        return (size_t)h;
    }
};

struct Node {   // IVoxNode<PointT, 3> (ivox3d_node.hpp:40-62): points in insertion order
    std::vector<std::array<float, 3>> pts;
    std::vector<int32_t> ids;
};

// nearby_grids_ (ivox3d.h:213-243): CENTER / NEARBY6 / NEARBY18 / NEARBY26 = the offsets with at most 0 / 1 / 2 / 3
// non-zero components (1, 7, 19, 27 voxels).  The visiting order does not influence the result set.
struct Deltas {
    int8_t d[27][3];
    Deltas() {
        int n = 0;
        for (int nz = 0; nz <= 3; ++nz)
            for (int x = -1; x <= 1; ++x)
                for (int y = -1; y <= 1; ++y)
                    for (int z = -1; z <= 1; ++z)
                        if ((x != 0) + (y != 0) + (z != 0) == nz) { d[n][0] = (int8_t)x; d[n][1] = (int8_t)y; d[n][2] = (int8_t)z; ++n; }
    }
};
const Deltas kDeltas;

inline float dist2_eigen(const float* a, const float* b) {   // distance2, ivox3d_node.hpp:13-16
    float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    return dx * dx + (dy * dy + dz * dz);
}

}  // namespace

struct orc_ivox {
    float resolution = 0.2f, inv_resolution = 5.f;
    int nearby = 19;
    size_t capacity = 1000000;
    int32_t next_id = 0;
    size_t evicted = 0;
    typedef std::list<std::pair<std::array<int32_t, 3>, Node>> Cache;
    Cache cache;                                                             // grids_cache_ (front = most recent)
    std::unordered_map<std::array<int32_t, 3>, Cache::iterator, KeyHash3> grids;   // grids_map_
    std::array<int32_t, 3> pos2grid(const float* p) const {
        return {(int32_t)roundf(p[0] * inv_resolution), (int32_t)roundf(p[1] * inv_resolution), (int32_t)roundf(p[2] * inv_resolution)};
    }
};

extern "C" orc_ivox* orc_ivox_create(float resolution, int nearby, size_t capacity) {
    orc_ivox* v = new orc_ivox();
    v->resolution = resolution;
    v->inv_resolution = (float)(1.0 / resolution);   // options_.inv_resolution_ = 1.0 / options_.resolution_ (ivox3d.h:63)
    v->nearby = nearby;
    if (capacity) v->capacity = capacity;
    return v;
}
extern "C" void orc_ivox_destroy(orc_ivox* v) { delete v; }

// IVox::AddPoints (ivox3d.h:247-276)
extern "C" void orc_ivox_add_points(orc_ivox* v, const float* xyz, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        const float* p = xyz + 3 * i;
        const int32_t pid = v->next_id++;
        auto key = v->pos2grid(p);
        auto it = v->grids.find(key);
        if (it == v->grids.end()) {
            v->cache.push_front({key, Node()});
            v->grids.insert({key, v->cache.begin()});
            v->cache.front().second.pts.push_back({p[0], p[1], p[2]});
            v->cache.front().second.ids.push_back(pid);
            if (v->grids.size() >= v->capacity) {   // (:262-265)
                v->grids.erase(v->cache.back().first);
                v->cache.pop_back();
                ++v->evicted;
            }
        } else {
            it->second->second.pts.push_back({p[0], p[1], p[2]});
            it->second->second.ids.push_back(pid);
            v->cache.splice(v->cache.begin(), v->cache, it->second);
            v->grids[key] = v->cache.begin();
        }
    }
}

extern "C" size_t orc_ivox_num_grids(orc_ivox* v) { return v->grids.size(); }   // NumValidGrids (:206-209)
extern "C" size_t orc_ivox_evicted(orc_ivox* v) { return v->evicted; }
extern "C" size_t orc_ivox_points(orc_ivox* v, float* xyz, int32_t* ids, size_t cap) {
    size_t n = 0;
    for (auto& g : v->cache)
        for (size_t j = 0; j < g.second.pts.size(); ++j) {
            if (n < cap) {
                if (xyz) { xyz[3 * n] = g.second.pts[j][0]; xyz[3 * n + 1] = g.second.pts[j][1]; xyz[3 * n + 2] = g.second.pts[j][2]; }
                if (ids) ids[n] = g.second.ids[j];
            }
            ++n;
        }
    return n;
}

// IVox::GetClosestPoint(pt, closest_pt, max_num, max_range) (ivox3d.h:132-204) over
// IVoxNode::KNNPointByCondition (ivox3d_node.hpp:141-206), batched.  out_count[q] = -1 marks "no candidate":
// the reference returns false there WITHOUT touching closest_pt (:168-170), so the caller's list keeps its
// previous content.
extern "C" void orc_ivox_knn(orc_ivox* v, const float* q_xyz, size_t nq, int k, double max_range,
                             int32_t* out_id, float* out_sqd, float* out_xyz, int32_t* out_count, int64_t* tie_straddles) {
    struct DP { float d; int32_t id; const float* p; };
    auto less = [](const DP& a, const DP& b) { return a.d < b.d || (a.d == b.d && a.id < b.id); };
    int64_t ties = 0;
    const double r2 = max_range * max_range;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : ties)
    for (long qi = 0; qi < (long)nq; ++qi) {
        const float* q = q_xyz + 3 * qi;
        const auto key = v->pos2grid(q);
        std::vector<DP> cand, all;
        for (int t = 0; t < v->nearby; ++t) {
            std::array<int32_t, 3> dk = {key[0] + kDeltas.d[t][0], key[1] + kDeltas.d[t][1], key[2] + kDeltas.d[t][2]};
            auto it = v->grids.find(dk);
            if (it == v->grids.end()) continue;
            const Node& nd = it->second->second;
            std::vector<DP> loc;
            for (size_t j = 0; j < nd.pts.size(); ++j) {
                double d = (double)dist2_eigen(nd.pts[j].data(), q);        // double d = distance2(pt, point) (:158)
                if (d < r2) loc.push_back({(float)d, nd.ids[j], nd.pts[j].data()});
            }
            all.insert(all.end(), loc.begin(), loc.end());
            if ((int)loc.size() > k) {                                       // nth_element + resize (:179-183)
                std::sort(loc.begin(), loc.end(), less);
                loc.resize(k);
            }
            cand.insert(cand.end(), loc.begin(), loc.end());
        }
        for (int s = 0; s < k; ++s) {
            out_id[qi * k + s] = -1; out_sqd[qi * k + s] = INFINITY;
            out_xyz[(qi * k + s) * 3] = out_xyz[(qi * k + s) * 3 + 1] = out_xyz[(qi * k + s) * 3 + 2] = 0.f;
        }
        if (cand.empty()) { out_count[qi] = -1; continue; }
        std::sort(cand.begin(), cand.end(), less);                           // two nth_element calls (:176-181): canonical order here
        const int found = std::min<int>(k, (int)cand.size());
        out_count[qi] = found;
        for (int s = 0; s < found; ++s) {
            out_id[qi * k + s] = cand[s].id; out_sqd[qi * k + s] = cand[s].d;
            for (int c = 0; c < 3; ++c) out_xyz[(qi * k + s) * 3 + c] = cand[s].p[c];
        }
        // a distance tie across rank k (over ALL candidates) makes the reference's choice implementation-defined
        std::sort(all.begin(), all.end(), less);
        if ((int)all.size() > k && all[k].d == all[k - 1].d) ties += 1;
    }
    if (tie_straddles) *tie_straddles = ties;
}

// map_incremental, iVox branch (src/laser_mapping.cpp:340-381): flag per point
//   1 = points_to_add, 2 = point_no_need_downsample, 0 = dropped.
// near_cnt[i] <= 0 stands for Nearest_Points[i].empty().  Eigen float vector arithmetic restated:
// center = ((pw / fs).floor() + 0.5) * fs with fs narrowed to float; calc_dist(Vector3f, Vector3f) = (p1 - p2).norm()
// (include/common_lib.h:85).
extern "C" void orc_ivox_map_filter(const float* world_xyz, size_t n, const float* near_xyz /* n*5*3 */, const int32_t* near_cnt,
                                    double filter_size_map_min, uint8_t* flag_out) {
    const float fsf = (float)filter_size_map_min;
    for (size_t i = 0; i < n; ++i) {
        const float* pw = world_xyz + 3 * i;
        int flag = 1;
        if (near_cnt[i] > 0) {
            float center[3];
            for (int c = 0; c < 3; ++c) center[c] = (floorf(pw[c] / fsf) + 0.5f) * fsf;
            const float* n0 = near_xyz + i * 15;
            const double half = 0.5 * filter_size_map_min;
            if ((double)fabsf(n0[0] - center[0]) > half && (double)fabsf(n0[1] - center[1]) > half && (double)fabsf(n0[2] - center[2]) > half) {
                flag = 2;
            } else {
                const float dist = sqrtf(dist2_eigen(pw, center));
                if (near_cnt[i] >= 5) {
                    for (int r = 0; r < 5; ++r)
                        if ((double)sqrtf(dist2_eigen(near_xyz + i * 15 + 3 * r, center)) < (double)dist + 1e-6) { flag = 0; break; }
                }
            }
        }
        flag_out[i] = (uint8_t)flag;
    }
}

// GNSSProcessing::computeH (src/GNSS_Processing.cpp:204-237), the block LaserMapping::h_share_model adds to HPH / HPL
// (src/laser_mapping.cpp:612-617).  gnss_pos is already in the global frame (:206-207).  rot row-major.
// HTH 6x6 row-major, HTL 6; both zero when |error| > 2 (:221-222).
extern "C" void orc_gnss_block(const double* rot, const double* pos, const double* gnss_pos, const double* antlever,
                               const double* std3, double* HTH36, double* HTL6) {
    for (int i = 0; i < 36; ++i) HTH36[i] = 0.0;
    for (int i = 0; i < 6; ++i) HTL6[i] = 0.0;
    const double* l = antlever;
    const double K[9] = {0.0, -l[2], l[1], l[2], 0.0, -l[0], -l[1], l[0], 0.0};   // SKEW_SYM_MATRX, include/so3_math.h:9
    double H[3][6], err[3];
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) {
            H[r][c] = rot[r * 3] * K[c] + rot[r * 3 + 1] * K[3 + c] + rot[r * 3 + 2] * K[6 + c];   // rot_end * [antlever]x
            H[r][3 + c] = (r == c) ? -1.0 : 0.0;
        }
        err[r] = gnss_pos[r] - pos[r] + (rot[r * 3] * l[0] + rot[r * 3 + 1] * l[1] + rot[r * 3 + 2] * l[2]);
    }
    if (sqrt(err[0] * err[0] + err[1] * err[1] + err[2] * err[2]) > 2.0) return;
    const double w[3] = {1.0 / std3[0], 1.0 / std3[1], 100.0 / std3[2]};
    for (int a = 0; a < 6; ++a) {
        for (int b = 0; b < 6; ++b)
            for (int k = 0; k < 3; ++k) HTH36[a * 6 + b] += H[k][a] * w[k] * H[k][b];
        for (int k = 0; k < 3; ++k) HTL6[a] += H[k][a] * w[k] * err[k];
    }
}
