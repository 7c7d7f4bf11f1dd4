/* ORACLE (test infrastructure).  The "checker API" exported identically by
 *   oracle/liboracle_port.so        -- CPU restatement (loam_port.hpp)
 *   oracle/liboracle_port_libm.so   -- same, glibc trig instead of detmath
 *   oracle/_ref/libloam_ref.so      -- the reference's own .cpp files compiled from /root/reference
 * One handle = the four nodes wired sr -> lo -> tm.onOdometry -> (lm -> tm.onMapping). */
#ifndef LOAM_CHECKER_API_H_
#define LOAM_CHECKER_API_H_
#include "loam_b200.h"
#ifdef __cplusplus
extern "C" {
#endif
const char* chk_kind(void);
void* chk_create(const loam_params* p);
void chk_destroy(void* h);
int chk_imu(void* h, double stamp, const double quat[4], const float acc[3]);
int chk_push_sweep(void* h, const float* xyz, int n, double stamp, loam_sweep_result* out);
int chk_get_cloud(void* h, int which, float* xyzi, int cap);
int chk_get_ints(void* h, int which, int* out, int cap);
int chk_get_floats(void* h, int which, float* out, int cap);
void chk_busy_seconds(void* h, double out[4]);
void chk_lo_set_state(void* h, const float transform[6], const float transformSum[6]);
/* the chain laserOdometry -> transformMaintenance -> laserMapping fed with HOST copies of scanRegistration's
 * six topics (x,y,z,intensity per point; imu_trans = the 12 floats of /imu_trans, NULL = zeros) */
int chk_lo_push(void* h, const float* full, int n_full, const float* sharp, int n_sharp, const float* less_sharp, int n_less_sharp,
                const float* flat, int n_flat, const float* less_flat, int n_less_flat, const float imu_trans[12], double stamp,
                loam_sweep_result* out);
/* port only: laserMapping alone on recorded laserOdometry output; "LOAMMAP1" map files */
int chk_lm_push(void* h, const float* corner, int nc, const float* surf, int ns, const float* full, int nf,
                const float transformSum[6], double stamp, loam_sweep_result* out);
int chk_lm_save_map(void* h, const char* path);
int chk_lm_load_map(void* h, const char* path);
/* port only: the solve step of one LM iteration (cv::solve, cv::eigen threshold scan, matP, projection) on
 * `batch` given normal equations, laserOdometry.cpp:765-797 (thr 10) / laserMapping.cpp:922-954 (thr 100) */
void chk_solve_step(const float* AtA, const float* AtB, int batch, float thr, int* isDegenerate, float* matP36, float* X6);
/* port and reference: the three nodes on three threads with depth-2 queues, as the reference's processes run */
double chk_run_pipelined(void* h, const float* xyz, const int* counts, const double* stamps, int n_sweeps, float* poses);
void chk_decode_packets(const unsigned char* packets, int n_packets, float* xyz);
/* port only: the packet front end (VLP-16 / HDL-64E by p->n_scans, ring_model) and a revolution of packets through the chain */
int chk_packet_front_end(const loam_params* p, const unsigned char* packets, int n_packets, float* xyz, int* ring, float* reltime);
int chk_push_packets(void* h, const unsigned char* packets, int n_packets, int packet_metadata, double stamp, loam_sweep_result* out);
#ifdef __cplusplus
}
#endif
#endif
