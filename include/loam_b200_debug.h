/* loam_b200_debug.h -- test and development hooks of libloam_b200.so (loam_debug_*).  NOT part of the
 * drop-in boundary (include/loam_b200.h): tests/ and scripts/ use them to check device building blocks in
 * isolation and to read timelines.  The clock64 phase stamps inside the persistent LM kernels exist only in
 * builds made with -DLOAM_PHASE_CLOCKS; in the default build loam_debug_phase_enable returns LOAM_ERR_ARG. */
#ifndef LOAM_B200_DEBUG_H_
#define LOAM_B200_DEBUG_H_

#include "loam_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- test hooks: device versions of the building blocks, host pointers in and out ---- */
/* fn: 0 sin 1 cos 2 atan 3 atan2(y,x) 4 asin 5 sqrt 6 x/y 7 x*y+x (unfused) -- loam/detmath.h on the device */
int loam_debug_math(int fn, const float* x, const float* y, float* out, int n);
/* op 0: QR solve (m x n, one rhs) 1: Jacobi eigen (values then row vectors) 2: LU inverse; `batch` systems */
int loam_debug_linalg(int op, const float* A, const float* b, float* out, int m, int n, int batch);
/* the solve step of one LM iteration (QR solve; at iteration 0: eigen threshold scan, matP, projection;
 * laserOdometry.cpp:765-797 / laserMapping.cpp:922-954) on `batch` normal equations: AtA [batch][36],
 * AtB [batch][6] -> is_degenerate [batch], matP [batch][36], X [batch][6].  mode 0 = as the LM kernels
 * run it (a positive-definiteness pre-test may skip the eigen-decomposition; matP is then the identity,
 * which is never read while is_degenerate is 0), mode 1 = always decompose, as the reference does */
int loam_debug_solve_step(const float* AtA, const float* AtB, int batch, float thr, int mode, int* is_degenerate,
                          float* matP, float* X);
/* matP (laserOdometry.cpp:405 / laserMapping.cpp:386) as of the node's last completed pass: node = a loam_lo*
 * (which 0) or a loam_lm* (which 1).  Only meaningful while the node's isDegenerate is set (see above) */
int loam_debug_matp(void* node, int which, float out36[36]);
/* the device packet decoder alone: xyz receives n_packets * 384 points (3 floats each, NaN = no return) */
int loam_debug_decode_packets(const unsigned char* packets, int n_packets, float* xyz);
/* the packet front end of the sensor `p` describes (16 rings / ring_model 0 = VLP-16, 64 rings / ring_model 1 =
 * HDL-64E): decoded xyz per slot and, when ring and reltime are given, ring id and relTime as
 * loam_params.packet_metadata = 1 derives them from the packet fields (ring -1 = no return) */
int loam_debug_packet_front_end(const loam_params* p, const unsigned char* packets, int n_packets, float* xyz, int* ring,
                                float* reltime);
int loam_debug_sort_pairs(unsigned int* keys, unsigned int* vals, int n);
int loam_debug_voxel_filter(const float* xyzi, int n, float leaf, float* out, int cap_out, int* n_out);
/* same filter through the one-kernel cooperative path laserMapping uses for the incoming sweep */
int loam_debug_voxel_filter_coop(const float* xyzi, int n, float leaf, float* out, int cap_out, int* n_out);

/* phase clocks (SM cycles, clock64 of CTA 0) of the persistent LM kernels: which 0 = laserOdometry,
 * 1 = laserMapping; slot 0 = number of stamps that follow */
int loam_debug_phase_enable(int on);
int loam_debug_phase_clocks(int which, long long* out, int n);

/* map-update statistics of the last mapped sweep, 2 x 8 ints (corner, surf): merge cubes, general-path
 * cubes, append-only cubes, max old points, max new points, points written, -, cubes gone dirty */
int loam_debug_lm_counters(loam_lm* h, int* out, int n);

/* timeline of the pipeline's streams: while enabled, every launch is bracketed by one-thread
 * kernels that write %globaltimer on the launch's stream (nothing is synchronised);
 * on = 3 stamps only the begin and end of every node pass (light enough not to disturb the
 * overlap).  loam_debug_trace_dump writes "kernel,stream,start_ns,end_ns" rows and returns their number */
int loam_debug_trace_enable(int on);
int loam_debug_trace_dump(const char* path);

/* nanoseconds the calling host threads spent blocked on the library's back-pressure events (staging
 * buffers, parameter slots) since the library was loaded: enqueue time minus this = launch cost */
long long loam_debug_host_wait_ns(void);
/* host nanoseconds the pipeline entry points spent enqueueing each node (sr, lo + record, lm) */
void loam_debug_host_node_ns(long long out[3]);
#ifdef __cplusplus
}
#endif
#endif /* LOAM_B200_DEBUG_H_ */
