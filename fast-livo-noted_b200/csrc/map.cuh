// map.cuh — host-side owner of the GPU-resident map (device tables + insert / delete pipelines).
// Replaces the internals of the reference's KD_TREE (include/ikd-Tree/ikd_Tree.{h,cpp}) while keeping
// its observable semantics; see flivo_dev.cuh for the layout and DESIGN.md for the parity argument.
#pragma once
#include "flivo_dev.cuh"
#include <string>

namespace flivo {

struct SegInfo {      // decision of one voxel segment of an Add_Points batch
    int keep_id;      // id of the stored point that survives (or -1)
    int new_idx;      // batch index of the point to store (or -1)
    int triggers;     // contribution to the reference's tmp_counter
    int flags;        // bit0: segment head, bit1: deferred to the serial pass, bit2: something changes
};

// per-call scalars of a captured insert sequence
struct BatchJob {
    unsigned int n;   // real points of the batch (the sequence runs on n rounded up to ADD_PAD items)
    int base_id;      // insertion id of the first point
};
constexpr size_t ADD_PAD = 8192;
constexpr size_t ADD_GRAPH_MAX = 1u << 18;   // batches up to this size are sized conservatively (no mid-sequence sync) and may be captured

class MapStore {
public:
    MapStore() = default;
    ~MapStore();
    // all calls return cudaError_t-like int (0 ok) and leave a message in err on failure
    int init(cudaStream_t stream, float ds, int cell_ratio, size_t initial_points, int ivox_nearby, std::string& err);
    void release();

    int set_downsample(float ds, std::string& err);
    // Build: replaces contents; points are device float4 with w ignored; ids 0..n-1
    int build(const float4* d_pts, size_t n, std::string& err);
    // Add_Points: ids = next_id .. next_id+n-1 (every offered point consumes one id)
    int add_points(const float4* d_pts, size_t n, bool downsample_on, int* n_added, std::string& err);
    int delete_boxes(const float* h_boxes6, size_t nb, int* n_deleted, std::string& err);
    int add_boxes(const float* h_boxes6, size_t nb, int* n_revived, std::string& err);
    int delete_points(const float4* d_pts, size_t n, int* n_deleted, std::string& err);
    int flatten(float* h_xyz, int* h_id, size_t cap, size_t* n, std::string& err);
    int range(float out6[6], std::string& err);
    int acquire_removed(float* h_xyz, size_t cap, size_t* n, std::string& err);
    int refresh_counters(std::string& err);   // device counters -> host mirror (synchronises the stream)

    const MapView& view() const { return v_; }
    const MapCounters& counters() const { return h_cnt_; }
    int64_t live() const { return (int64_t)h_cnt_.live; }
    float ds() const { return v_.ds; }
    size_t dir_cap() const { return dir_cap_; }
    size_t bucket_cap() const { return bucket_cap_; }
    bool empty() const { return h_cnt_.live == 0; }
    int next_id() const { return next_id_; }
    // the waits of Add_Points sleep on an event (cudaEventBlockingSync) instead of spinning in cudaStreamSynchronize: for
    // callers that drive many contexts from more host threads than they have cores (flivo_frame, see api.cu)
    void set_blocking_wait(bool on) { block_wait_ = on; }

private:
    int alloc_tables(size_t dir_cap, size_t bucket_cap, std::string& err);
    int ensure_capacity(size_t new_blocks_ub, size_t new_buckets_ub, std::string& err);
    int grow_dir(size_t new_cap, std::string& err);
    int grow_buckets(size_t new_cap, std::string& err);
    int ensure_scratch(size_t n, std::string& err);
    // common pipeline; d_pts_id: float4 with w = id bits
    int insert_batch(const float4* d_pts_id, size_t n, bool downsample_on, int* n_added, std::string& err);
    int stamp_ids(const float4* d_pts, size_t n, int base_id, std::string& err);  // -> scratch_pts_
    // the kernels of one insert batch, k_point_keys .. k_block_records, for n items; with a job block the items at and
    // beyond job->n are padding (invalid keys, skipped by every stage)
    int enqueue_insert(const float4* d_pts_id, size_t n, bool downsample_on, const BatchJob* job, bool keys_sorted, std::string& err);
    int add_points_graph(const float4* d_pts, size_t n, int* n_added, std::string& err);
    int finish_counters(std::string& err);   // host mirror <- pinned copy of the device counters
    void drop_add_graph();

    cudaStream_t stream_ = nullptr;
    MapView v_{};
    MapCounters* d_cnt_ = nullptr;
    MapCounters h_cnt_{};
    MapCounters* h_pinned_ = nullptr;
    size_t dir_cap_ = 0, bucket_cap_ = 0;
    int next_id_ = 0;
    int cell_ratio_ = 2;
    size_t blocks_ub_ = 0, buckets_ub_ = 0;   // host-side upper bounds between counter refreshes
    // removed-points log (Points_deleted of the reference, ikd_Tree.h:143)
    float4* d_removed_ = nullptr;
    size_t removed_cap_ = 0;
    // scratch for insert batches
    size_t scratch_n_ = 0;
    float4* scratch_pts_ = nullptr;
    unsigned long long *keys_a_ = nullptr, *keys_b_ = nullptr;
    unsigned int *idx_a_ = nullptr, *idx_b_ = nullptr, *defer_list_ = nullptr;
    unsigned char* defer_flag_ = nullptr;
    SegInfo* seg_ = nullptr;
    float4* slop_list_ = nullptr;   // batch points whose voxel box membership is ambiguous (x,y,z,batch idx)
    void* cub_tmp_ = nullptr;
    size_t cub_tmp_bytes_ = 0;
    unsigned int* d_misc_ = nullptr;  // [0]=unique blocks, [1]=unique cells, [2]=slop count, [3]=deferred points
    // Add_Points(…, true) of a scan-sized batch as ONE captured graph (map_incremental is ~30 launches otherwise and
    // launch-bound): per-call scalars travel in a BatchJob block, the graph depends on the padded size, the source
    // pointer and the map's tables only
    BatchJob* d_bjob_ = nullptr;
    BatchJob* h_bjob_ = nullptr;      // pinned
    cudaGraphExec_t add_exec_ = nullptr;
    const float4* add_src_ = nullptr;
    size_t add_npad_ = 0;
    MapView add_view_{};
    const void* add_scratch_ = nullptr;
    int add_graph_fail_ = 0;
    bool block_wait_ = false;
    cudaEvent_t block_ev_ = nullptr;
    int wait_stream(std::string& err);   // cudaStreamSynchronize(stream_), or the blocking event wait
};

constexpr int SLOP_LIST_CAP = 4096;

}  // namespace flivo
