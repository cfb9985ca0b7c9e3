// frontend.cuh — device-side front end of the path: motion compensation and voxel-grid down-sampling of the raw scan
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace flivo {

constexpr int UNDISTORT_MAX_POSES = 256;   // IMU samples of one scan (+1); 200 Hz x 0.1 s = 21 in the reference's data

// msg/Pose6D.msg, filled by set_pose6d (include/common_lib.h:618-633)
struct PoseDev {
    double offset_time;
    double acc[3], gyr[3], vel[3], pos[3], rot[9];
};

struct UndistortParams {
    const PoseDev* poses;   // device, n_poses entries, offset_time ascending
    int n_poses;
    double pos_end[3];      // state_inout.pos_end after the forward propagation
    double R_LI[9], t_LI[3];
    double extR_Ri[9];      // Lid_rot_to_IMU^T * rot_end^T        (src/IMU_Processing.cpp:336)
    double exrR_extT[3];    // Lid_rot_to_IMU^T * Lid_offset_to_IMU (:337)
};

struct VoxelGridState {
    unsigned min_ord[3], max_ord[3];   // order-preserving images of the float bounds
    unsigned n_finite;
    int min_b[3], div_b[3], mul[3];
    int overflow;                      // leaf size too small for the extent (index would overflow int)
    unsigned n_out;
};

void launch_undistort(float4* pts, const float* t_ms, size_t n, const UndistortParams& U, cudaStream_t stream);
size_t voxel_grid_temp_bytes(size_t cap);
cudaError_t launch_voxel_grid(const float4* pts, const float* t_ms, size_t n, float leaf, VoxelGridState* st, unsigned* keys_a, unsigned* keys_b,
                              unsigned* idx_a, unsigned* idx_b, unsigned* head, unsigned* pos, void* tmp, size_t tmp_bytes, float4* out, float* out_t,
                              cudaStream_t stream);
cudaError_t launch_voxel_passthrough(const float4* pts, const float* t_ms, size_t n, VoxelGridState* st, float4* out, float* out_t, cudaStream_t stream);

}  // namespace flivo
