// imu_processing.h — ROS-free drop-in for the part of the reference's ImuProcess that feeds the LIO path
// (include/IMU_Processing.h:44-98, src/IMU_Processing.cpp:200-376 `UndistortPcl`), LiDAR-only frames
// (lidar_meas.is_lidar_end == true, no camera sub-frames).
//
// The forward propagation of the state and its 18x18 covariance over the IMU samples of the frame is a short
// sequential recurrence (20-odd samples): it runs on the host exactly as in the reference (:237-334).  The per-point
// back-propagation (:336-376), the loop over 10^5..10^6 raw points, runs on the device: the raw scan is copied to HBM
// once (flivo_raw_upload) and stays there for the voxel-grid down-sampling and the measurement update that follow.
// Header-only; link with -lflivo_b200.
#pragma once
#include <math.h>
#include <string.h>
#include <deque>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/flivo.h"
#include "flivo_types.h"

// sensor_msgs::Imu, the three fields UndistortPcl reads
struct ImuMsg {
    double stamp = 0.0;                      // header.stamp.toSec()
    double angular_velocity[3] = {0, 0, 0};
    double linear_acceleration[3] = {0, 0, 0};
};

// MeasureGroup / LidarMeasureGroup (include/common_lib.h) reduced to what a LiDAR-only frame carries
struct LidarMeasureGroup {
    double lidar_beg_time = 0.0;
    double last_update_time = 0.0;
    std::deque<ImuMsg> imu;                  // measures.back().imu
    PointVector lidar;                       // lidar->points: curvature = time offset in ms, ascending
};

class ImuProcess {
public:
    static constexpr double G_m_s2 = 9.81;   // include/common_lib.h:31
    ImuProcess() { Reset(); }

    void Reset() {                           // src/IMU_Processing.cpp:36-49
        set3(mean_acc, 0, 0, -1.0);
        set3(angvel_last, 0, 0, 0); set3(acc_s_last, 0, 0, 0);
        start_timestamp_ = -1; last_lidar_end_time_ = 0.0;
        last_imu_ = ImuMsg();
        IMUpose.clear();
    }
    void set_extrinsic(const double transl[3], const double rot[9]) {    // :66-70
        memcpy(Lid_offset_to_IMU, transl, 24); memcpy(Lid_rot_to_IMU, rot, 72);
    }
    void set_gyr_bias_cov(const double b[3]) { memcpy(cov_bias_gyr, b, 24); }
    void set_acc_bias_cov(const double b[3]) { memcpy(cov_bias_acc, b, 24); }

    // src/IMU_Processing.cpp:200-376.  The raw scan of lidar_meas is uploaded to `ctx` and compensated there
    // (-> feats_undistort, resident in HBM); state_inout receives the state and covariance propagated to the scan end.
    void UndistortPcl(flivo_ctx* ctx, LidarMeasureGroup& lidar_meas, StatesGroup& state_inout) {
        std::deque<ImuMsg> v_imu = lidar_meas.imu;                                       // :207-208
        v_imu.push_front(last_imu_);
        const double imu_end_time = v_imu.back().stamp;
        const double pcl_beg_time = lidar_meas.lidar_beg_time > lidar_meas.last_update_time ? lidar_meas.lidar_beg_time : lidar_meas.last_update_time;   // :211
        start_timestamp_ = pcl_beg_time;
        if (lidar_meas.lidar.empty()) throw std::runtime_error("ImuProcess::UndistortPcl: empty scan");
        const double pcl_end_time = lidar_meas.lidar_beg_time + lidar_meas.lidar.back().curvature / double(1000);   // :218-220
        lidar_meas.last_update_time = pcl_end_time;                                      // :230

        IMUpose.clear();                                                                 // :237-238
        IMUpose.push_back(set_pose6d(0.0, acc_s_last, angvel_last, state_inout.vel_end, state_inout.pos_end, state_inout.rot_end));
        double acc_imu[3], angvel_avr[3], acc_avr[3], vel_imu[3], pos_imu[3], R_imu[9];
        memcpy(acc_imu, acc_s_last, 24); memcpy(angvel_avr, angvel_last, 24);
        memcpy(vel_imu, state_inout.vel_end, 24); memcpy(pos_imu, state_inout.pos_end, 24); memcpy(R_imu, state_inout.rot_end, 72);
        const double mean_acc_norm = sqrt(mean_acc[0] * mean_acc[0] + mean_acc[1] * mean_acc[1] + mean_acc[2] * mean_acc[2]);
        double dt = 0;
        for (size_t k = 0; k + 1 < v_imu.size(); ++k) {                                  // :246
            const ImuMsg& head = v_imu[k];
            const ImuMsg& tail = v_imu[k + 1];
            if (tail.stamp < last_lidar_end_time_) continue;                             // :251
            for (int a = 0; a < 3; ++a) {
                angvel_avr[a] = 0.5 * (head.angular_velocity[a] + tail.angular_velocity[a]);
                acc_avr[a] = 0.5 * (head.linear_acceleration[a] + tail.linear_acceleration[a]);
            }
            for (int a = 0; a < 3; ++a) {                                                // :266-267
                angvel_avr[a] -= state_inout.bias_g[a];
                acc_avr[a] = acc_avr[a] * G_m_s2 / mean_acc_norm - state_inout.bias_a[a];
            }
            dt = (head.stamp < last_lidar_end_time_) ? tail.stamp - last_lidar_end_time_ : tail.stamp - head.stamp;   // :269-276
            propagate_cov(state_inout, R_imu, angvel_avr, acc_avr, dt);                  // :278-298
            double Exp_f[9], Rn[9];
            exp_rate(angvel_avr, dt, Exp_f);
            mm3(R_imu, Exp_f, Rn); memcpy(R_imu, Rn, 72);                                // :301
            for (int a = 0; a < 3; ++a)                                                  // :304
                acc_imu[a] = (R_imu[a * 3] * acc_avr[0] + R_imu[a * 3 + 1] * acc_avr[1] + R_imu[a * 3 + 2] * acc_avr[2]) + state_inout.gravity[a];
            for (int a = 0; a < 3; ++a) pos_imu[a] = (pos_imu[a] + vel_imu[a] * dt) + 0.5 * acc_imu[a] * dt * dt;   // :307
            for (int a = 0; a < 3; ++a) vel_imu[a] = vel_imu[a] + acc_imu[a] * dt;      // :310
            memcpy(angvel_last, angvel_avr, 24); memcpy(acc_s_last, acc_imu, 24);        // :313-314
            IMUpose.push_back(set_pose6d(tail.stamp - pcl_beg_time, acc_imu, angvel_avr, vel_imu, pos_imu, R_imu));   // :315-316
        }
        {   // pose at the frame end (:320-335)
            const bool from_imu_end = imu_end_time > pcl_beg_time;
            const double ref_t = from_imu_end ? imu_end_time : pcl_beg_time;
            const double note = pcl_end_time > ref_t ? 1.0 : -1.0;
            dt = note * (pcl_end_time - ref_t);
            double w[3] = {note * angvel_avr[0], note * angvel_avr[1], note * angvel_avr[2]}, E[9];
            exp_rate(w, dt, E);
            for (int a = 0; a < 3; ++a) state_inout.vel_end[a] = vel_imu[a] + note * acc_imu[a] * dt;
            mm3(R_imu, E, state_inout.rot_end);
            for (int a = 0; a < 3; ++a) state_inout.pos_end[a] = (pos_imu[a] + note * vel_imu[a] * dt) + note * 0.5 * acc_imu[a] * dt * dt;
        }
        last_imu_ = v_imu.back();                                                        // :337-338
        last_lidar_end_time_ = pcl_end_time;

        // :336-376 on the device
        ck(ctx, flivo_raw_upload(ctx, &lidar_meas.lidar[0].x, lidar_meas.lidar.size(), sizeof(PointType), 36, 32), "raw upload");
        if (IMUpose.size() >= 2)
            ck(ctx, flivo_raw_undistort(ctx, IMUpose.data(), (int)IMUpose.size(), state_inout.rot_end, state_inout.pos_end, Lid_rot_to_IMU, Lid_offset_to_IMU), "undistort");
    }

    // feats_undistort back on the host (only when a caller wants to look at it: the pipeline itself keeps it in HBM)
    void download_undistorted(flivo_ctx* ctx, PointVector& pcl_out, const LidarMeasureGroup& lidar_meas) {
        size_t n = 0;
        ck(ctx, flivo_raw_download(ctx, nullptr, 0, &n), "raw size");
        std::vector<float> buf(n * 4);
        if (n) ck(ctx, flivo_raw_download(ctx, buf.data(), n, &n), "raw download");
        pcl_out.assign(lidar_meas.lidar.begin(), lidar_meas.lidar.begin() + n);
        for (size_t i = 0; i < n; ++i) { pcl_out[i].x = buf[4 * i]; pcl_out[i].y = buf[4 * i + 1]; pcl_out[i].z = buf[4 * i + 2]; }
    }

    std::vector<flivo_pose6d> IMUpose;
    double cov_acc[3] = {0.1, 0.1, 0.1}, cov_gyr[3] = {0.1, 0.1, 0.1};                   // :22-27
    double cov_bias_gyr[3] = {0.1, 0.1, 0.1}, cov_bias_acc[3] = {0.1, 0.1, 0.1};
    double mean_acc[3];
    double angvel_last[3], acc_s_last[3];
    double Lid_offset_to_IMU[3] = {0, 0, 0};
    double Lid_rot_to_IMU[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    double start_timestamp_, last_lidar_end_time_;
    ImuMsg last_imu_;

    // Exp(ang_vel, dt), include/so3_math.h:30-52
    static void exp_rate(const double w[3], double dt, double R[9]) {
        const double nrm = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
        for (int e = 0; e < 9; ++e) R[e] = (e % 4 == 0) ? 1.0 : 0.0;
        if (nrm > 0.0000001) {
            const double a[3] = {w[0] / nrm, w[1] / nrm, w[2] / nrm};
            const double K[9] = {0.0, -a[2], a[1], a[2], 0.0, -a[0], -a[1], a[0], 0.0};
            const double ang = nrm * dt, s = sin(ang), c1 = 1.0 - cos(ang);
            double cK[9], KK[9];
            for (int e = 0; e < 9; ++e) cK[e] = c1 * K[e];
            mm3(cK, K, KK);
            for (int e = 0; e < 9; ++e) R[e] = (R[e] + s * K[e]) + KK[e];
        }
    }

private:
    static void set3(double* v, double a, double b, double c) { v[0] = a; v[1] = b; v[2] = c; }
    static void mm3(const double* A, const double* B, double* C) {
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) C[i * 3 + j] = A[i * 3] * B[j] + A[i * 3 + 1] * B[3 + j] + A[i * 3 + 2] * B[6 + j];
    }
    static flivo_pose6d set_pose6d(double t, const double a[3], const double g[3], const double v[3], const double p[3], const double R[9]) {
        flivo_pose6d k;                                                                  // include/common_lib.h:618-633
        k.offset_time = t;
        memcpy(k.acc, a, 24); memcpy(k.gyr, g, 24); memcpy(k.vel, v, 24); memcpy(k.pos, p, 24); memcpy(k.rot, R, 72);
        return k;
    }
    // state.cov = F_x * state.cov * F_x^T + cov_w   (:278-298), 18x18, order rot, pos, vel, bg, ba, grav
    void propagate_cov(StatesGroup& st, const double R_imu[9], const double angvel_avr[3], const double acc_avr[3], double dt) const {
        const int N = 18;
        std::vector<double> F(N * N, 0.0), W(N * N, 0.0), T(N * N), O(N * N);
        for (int i = 0; i < N; ++i) F[i * N + i] = 1.0;
        double Em[9];
        exp_rate(angvel_avr, -dt, Em);
        const double sk[9] = {0.0, -acc_avr[2], acc_avr[1], acc_avr[2], 0.0, -acc_avr[0], -acc_avr[1], acc_avr[0], 0.0};
        double Rs[9];
        mm3(R_imu, sk, Rs);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) {
                F[i * N + j] = Em[i * 3 + j];                                   // block(0,0)
                F[i * N + 9 + j] = (i == j) ? -dt : 0.0;                        // block(0,9)  = -I dt
                F[(3 + i) * N + 6 + j] = (i == j) ? dt : 0.0;                   // block(3,6)  =  I dt
                F[(6 + i) * N + j] = -Rs[i * 3 + j] * dt;                       // block(6,0)  = -R [a]x dt
                F[(6 + i) * N + 12 + j] = -R_imu[i * 3 + j] * dt;               // block(6,12) = -R dt
                F[(6 + i) * N + 15 + j] = (i == j) ? dt : 0.0;                  // block(6,15) =  I dt
            }
        for (int i = 0; i < 3; ++i) {
            W[i * N + i] = cov_gyr[i] * dt * dt;
            W[(9 + i) * N + 9 + i] = cov_bias_gyr[i] * dt * dt;
            W[(12 + i) * N + 12 + i] = cov_bias_acc[i] * dt * dt;
            for (int j = 0; j < 3; ++j) {                                       // R diag(cov_acc) R^T dt^2
                double s = 0.0;
                for (int k = 0; k < 3; ++k) s += R_imu[i * 3 + k] * cov_acc[k] * R_imu[j * 3 + k];
                W[(6 + i) * N + 6 + j] = s * dt * dt;
            }
        }
        for (int i = 0; i < N; ++i)
            for (int j = 0; j < N; ++j) {
                double s = 0.0;
                for (int k = 0; k < N; ++k) s += F[i * N + k] * st.cov[k * N + j];
                T[i * N + j] = s;
            }
        for (int i = 0; i < N; ++i)
            for (int j = 0; j < N; ++j) {
                double s = 0.0;
                for (int k = 0; k < N; ++k) s += T[i * N + k] * F[j * N + k];
                O[i * N + j] = s + W[i * N + j];
            }
        memcpy(st.cov, O.data(), sizeof(double) * N * N);
    }
    static void ck(flivo_ctx* ctx, int rc, const char* what) {
        if (rc != FLIVO_OK) throw std::runtime_error(std::string("ImuProcess::") + what + ": " + flivo_last_error(ctx));
    }
};
