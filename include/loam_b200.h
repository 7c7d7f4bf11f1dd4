/* loam_b200.h -- C ABI of the B200-native LOAM per-sweep pipeline.
 *
 * Drop-in boundary: the reference (Maofei/loam_velodyne-1) has no linkable
 * API; its contract is the ROS topic graph between four nodes.  Each entry point below replaces
 * one ROS callback / loop body of the reference and keeps its argument meaning, ordering rules
 * and "silently wait for more data" behaviour.  All functions return 0 (LOAM_OK), 1
 * (LOAM_NOT_READY: the reference would `return`/`continue` without publishing) or a negative
 * error; no C++ exception crosses this boundary; no torch / CUDA types appear in signatures.
 * A handle is bound to one CUDA device and one stream and is NOT thread-safe; distinct handles
 * are independent.  Point clouds are device-resident; use the *_download calls to read them.
 * There is no CPU fallback: every call fails with LOAM_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef LOAM_B200_H_
#define LOAM_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LOAM_OK 0
#define LOAM_NOT_READY 1
#define LOAM_ERR_ARG (-1)
#define LOAM_ERR_CUDA (-2)
#define LOAM_ERR_CAPACITY (-3)

/* Every tunable that is a compile-time literal in the reference and that BASELINE.json's
 * configs need to vary.  All other constants are fixed to the reference's. */
typedef struct loam_params {
  int n_scans;          /* scanRegistration.cpp:61  N_SCANS (16)                                  */
  int ring_model;       /* 0: VLP-16 1-degree rounding, scanRegistration.cpp:248-256               */
                        /* 1: linear lower..upper, the HDL-64E block of the bk variant (:268-274)  */
  float ring_lower_deg; /* ring_model 1: elevation of ring 0                                      */
  float ring_upper_deg; /* ring_model 1: elevation of ring n_scans-1                              */
  int max_points;       /* scanRegistration.cpp:63-66, laserOdometry.cpp:84-91 array size (40000)  */
  int system_delay;     /* scanRegistration.cpp:57 (20); benchmarks use 0 on both sides           */
  int lo_max_iter;      /* laserOdometry.cpp:470 (25)                                             */
  int lm_max_iter;      /* laserMapping.cpp:710 (10)                                              */
  int map_arena_points; /* product only: capacity of each cube-map arena, in points               */
  int coop_ctas;        /* product only: CTAs of the persistent cooperative LM kernels (0 = one per SM); */
                        /* smaller grids let several handles' kernels run side by side (configs[4])   */
  int dev_flags;        /* product only: development A/B switches (0 = the shipped configuration)            */
  int packet_metadata;  /* raw-packet front end (loam_*_packets): 0 = decode the packets to xyz and run the     */
                        /* reference's SR-1 unchanged (ring from atan, relTime from atan2 + halfPassed,          */
                        /* scanRegistration.cpp:248-284); 1 = ring id from the laser id and relTime from the     */
                        /* packet's azimuth field, fused with the decoder: no atan / atan2 per point, no        */
                        /* halfPassed scan (csrc/pk_decode.cuh states the arithmetic and   */
                        /* its difference to the reference)                                                     */
  int reserved[4];
} loam_params;

/* Hard limits of this build (compile-time constants of the kernels; the reference has none of them because it
 * sizes std::vectors on the fly).  A sweep or a parameter block that exceeds one is refused with LOAM_ERR_ARG /
 * LOAM_ERR_CAPACITY and a loam_last_error() text that names the limit; nothing is truncated silently.
 *   LOAM_LIMIT_RINGS            n_scans                                   (SR_MAX_RINGS,  csrc/sr_kernels.cuh)
 *   LOAM_LIMIT_RING_POINTS      less-flat candidates of one ring           (SR_RING_CAP:   shared-memory voxel filter)
 *   LOAM_LIMIT_SECTOR_POINTS    points in one sixth of a ring              (SR_SECTOR_CAP: shared-memory sort)
 *   LOAM_LIMIT_CUBES_PER_SWEEP  cubes one sweep inserts into, per type     (PROC_MAX,      csrc/lm_kernels.cuh)
 *   LOAM_LIMIT_LO_ITERATIONS    lo_max_iter                                (one lane per accumulated row slab)
 * BASELINE.json's largest sensor (128 rings x 2048 azimuths) uses half of the per-ring limits. */
#define LOAM_LIMIT_RINGS 128
#define LOAM_LIMIT_RING_POINTS 4096
#define LOAM_LIMIT_SECTOR_POINTS 1024
#define LOAM_LIMIT_CUBES_PER_SWEEP 250
#define LOAM_LIMIT_LO_ITERATIONS 32

/* preset: 1 = VLP-16 (reference defaults), 2 = HDL-64E, 4 = 128-ring (BASELINE.json configs) */
int loam_params_preset(loam_params* p, int preset);

/* cloud selectors for the download calls */
enum loam_cloud {
  LOAM_CLOUD_FULL = 0,        /* /velodyne_cloud_2        scanRegistration.cpp:584-588 */
  LOAM_CLOUD_SHARP = 1,       /* /laser_cloud_sharp       :590-594 */
  LOAM_CLOUD_LESS_SHARP = 2,  /* /laser_cloud_less_sharp  :596-600 */
  LOAM_CLOUD_FLAT = 3,        /* /laser_cloud_flat        :602-606 */
  LOAM_CLOUD_LESS_FLAT = 4,   /* /laser_cloud_less_flat   :608-612 */
  LOAM_CLOUD_CORNER_LAST = 5, /* /laser_cloud_corner_last laserOdometry.cpp:913-917 */
  LOAM_CLOUD_SURF_LAST = 6,   /* /laser_cloud_surf_last   :919-923 */
  LOAM_CLOUD_FULL_END = 7,    /* /velodyne_cloud_3        :925-929 */
  LOAM_CLOUD_REGISTERED = 8,  /* /velodyne_cloud_registered laserMapping.cpp:1060-1069 */
  LOAM_CLOUD_SURROUND = 9,    /* /laser_cloud_surround    :1042-1057 */
  LOAM_CLOUD_MAP_CORNER = 10, /* laserCloudCornerFromMap  :674-679 (local map used by the last LM) */
  LOAM_CLOUD_MAP_SURF = 11,   /* laserCloudSurfFromMap */
  LOAM_CLOUD_STACK_CORNER = 12, /* laserCloudCornerStack  :693-696 */
  LOAM_CLOUD_STACK_SURF = 13    /* laserCloudSurfStack    :698-701 */
};

/* int-array selectors (debug / parity) */
enum loam_ints {
  LOAM_INTS_LABEL = 0,       /* cloudLabel[]           scanRegistration.cpp:66  */
  LOAM_INTS_PICKED = 1,      /* cloudNeighborPicked[]  :65 (state after selection) */
  LOAM_INTS_SORT_IND = 2,    /* cloudSortInd[]         :64 */
  LOAM_INTS_SHARP_IDX = 3,   /* indices (into FULL) of the sharp cloud, emission order */
  LOAM_INTS_LESS_SHARP_IDX = 4,
  LOAM_INTS_FLAT_IDX = 5,
  LOAM_INTS_CORNER_IND1 = 6, /* pointSearchCornerInd1[] laserOdometry.cpp:85 (last refresh) */
  LOAM_INTS_CORNER_IND2 = 7,
  LOAM_INTS_SURF_IND1 = 8,
  LOAM_INTS_SURF_IND2 = 9,
  LOAM_INTS_SURF_IND3 = 10,
  LOAM_INTS_SCAN_START = 11, /* scanStartInd[] scanRegistration.cpp:221 */
  LOAM_INTS_SCAN_END = 12
};
enum loam_floats {
  LOAM_FLOATS_CURVATURE = 0  /* cloudCurvature[] scanRegistration.cpp:63 */
};

/* per-sweep result of the whole pipeline */
typedef struct loam_sweep_result {
  int sr_status, lo_status, lm_status; /* LOAM_OK / LOAM_NOT_READY per node */
  int n_full, n_sharp, n_less_sharp, n_flat, n_less_flat;
  int lo_iters;         /* LM iterations executed by laserOdometry (0..25) */
  int lo_rows;          /* rows in the last odometry normal equations */
  int lo_degenerate;    /* isDegenerate, laserOdometry.cpp:404 */
  int lm_iters, lm_rows, lm_degenerate;
  int n_map_corner, n_map_surf, n_stack_corner, n_stack_surf;
  float transform[6];           /* laserOdometry.cpp:93  sweep-to-sweep motion */
  float transform_sum[6];       /* laserOdometry.cpp:94  /laser_odom_to_init */
  float transform_aft_mapped[6];/* laserMapping.cpp:100  /aft_mapped_to_init pose */
  float transform_bef_mapped[6];/* laserMapping.cpp:99   (twist fields of that message) */
  float transform_mapped[6];    /* transformMaintenance.cpp:51 /integrated_to_init */
  float imu_trans[12];          /* scanRegistration.cpp:614-629 */
} loam_sweep_result;

/* ---- scanRegistration node (scanRegistration.cpp) ---- */
typedef struct loam_sr loam_sr;
int loam_sr_create(const loam_params* p, int device, loam_sr** out);
void loam_sr_destroy(loam_sr* h);
/* imuHandler, scanRegistration.cpp:638-660.  quat = (x,y,z,w) orientation, acc = linear_acceleration */
int loam_sr_imu(loam_sr* h, double stamp, const double quat[4], const float acc[3]);
/* laserCloudHandler, :211-636.  xyz: n points, `stride_bytes` apart, 3 floats each, HOST memory,
 * sensor frame (x fwd, y left, z up), firing order.  NOT_READY while systemDelay swallows sweeps. */
int loam_sr_process(loam_sr* h, const float* xyz, int n, int stride_bytes, double stamp);
/* raw-packet front end: the same handler fed with the Velodyne VLP-16 data packets of
 * one revolution (n_packets x 1206 bytes, HOST memory, single-return mode) instead of the decoded
 * /velodyne_points cloud that scanRegistration.cpp:675 subscribes to.  The packets are decoded on
 * the device (csrc/pk_decode.cuh states the arithmetic: the published packet format with an ideal
 * calibration, azimuth interpolated inside a block pair as the velodyne_pointcloud driver does);
 * no-return slots become NaN and are dropped by the node's NaN filter (:228). */
#define LOAM_PACKET_BYTES 1206             /* VLP-16 and HDL-64E data packets alike */
#define LOAM_POINTS_PER_PACKET 384
#define LOAM_VLP16_PACKET_BYTES 1206
#define LOAM_VLP16_POINTS_PER_PACKET 384
int loam_sr_process_packets(loam_sr* h, const unsigned char* packets, int n_packets, double stamp);
int loam_sr_download(loam_sr* h, int which_cloud, float* xyzi, int cap_points, int* n);
int loam_sr_download_ints(loam_sr* h, int which, int* out, int cap, int* n);
int loam_sr_download_floats(loam_sr* h, int which, float* out, int cap, int* n);
int loam_sr_imu_trans(loam_sr* h, float out12[12]);

/* ---- laserOdometry node (laserOdometry.cpp main loop :410-931) ---- */
typedef struct loam_lo loam_lo;
int loam_lo_create(const loam_params* p, int device, loam_lo** out);
void loam_lo_destroy(loam_lo* h);
/* consumes the six topics scanRegistration just produced (same device).  NOT_READY on the first
 * sweep (:427-456).  *has_mapping_input = 1 on the sweeps whose clouds the reference publishes
 * to laserMapping (every (skipFrameNum+1)-th, :910-930). */
int loam_lo_process(loam_lo* h, loam_sr* features, int* has_mapping_input);
/* the same loop-body pass fed from HOST copies of the six topics (clouds as x,y,z,intensity per point,
 * ring-major as scanRegistration emits them; imu_trans = the 12 floats of /imu_trans, NULL = zeros):
 * the form for features that were computed, recorded or transported elsewhere */
int loam_lo_process_host(loam_lo* h, const float* full, int n_full, const float* sharp, int n_sharp,
                         const float* less_sharp, int n_less_sharp, const float* flat, int n_flat,
                         const float* less_flat, int n_less_flat, const float imu_trans[12], int* has_mapping_input);
int loam_lo_download(loam_lo* h, int which_cloud, float* xyzi, int cap_points, int* n);
int loam_lo_download_ints(loam_lo* h, int which, int* out, int cap, int* n);
int loam_lo_state(loam_lo* h, float transform[6], float transform_sum[6], int stats[4]);
/* teacher forcing for parity tests: overwrite transform / transformSum (laserOdometry.cpp:93-94) */
int loam_lo_set_state(loam_lo* h, const float transform[6], const float transform_sum[6]);

/* ---- laserMapping node (laserMapping.cpp main loop :408-1096) ---- */
typedef struct loam_lm loam_lm;
int loam_lm_create(const loam_params* p, int device, loam_lm** out);
void loam_lm_destroy(loam_lm* h);
/* imuHandler, laserMapping.cpp:323-335 */
int loam_lm_imu(loam_lm* h, double stamp, const double quat[4]);
/* consumes corner_last / surf_last / velodyne_cloud_3 / laser_odom_to_init of `odom` */
int loam_lm_process(loam_lm* h, loam_lo* odom, double stamp);
/* the same loop-body pass fed from HOST copies of the four topics (x,y,z,intensity per point;
 * transform_sum = laserOdometry's transformSum before the message encoding) -- the form a caller
 * that records / replays laserOdometry's output uses (BASELINE.json configs[2], mapping only) */
int loam_lm_process_host(loam_lm* h, const float* corner_last, int n_corner, const float* surf_last, int n_surf,
                         const float* full_res, int n_full, const float transform_sum[6], double stamp);
/* the form a ROS node uses (adapters/ros/): transform_sum_decoded is what laserOdometryHandler
 * (laserMapping.cpp:304-321) obtained from the /laser_odom_to_init message with tf getRPY, so the
 * library must not apply the message round trip again */
int loam_lm_process_msg(loam_lm* h, const float* corner_last, int n_corner, const float* surf_last, int n_surf,
                        const float* full_res, int n_full, const float transform_sum_decoded[6], double stamp);
/* asynchronous form with DEVICE-resident clouds (float4 x,y,z,intensity), enqueue only: nothing is
 * copied but the odometry state; the arrays must stay unchanged until loam_lm_sync() (which also
 * refreshes what loam_lm_state / loam_lm_download report).  loam_lm_stream() is the CUDA stream
 * the node's passes are ordered on. */
int loam_lm_process_device(loam_lm* h, const float* d_corner_last, int n_corner, const float* d_surf_last, int n_surf,
                           const float* d_full_res, int n_full, const float transform_sum[6], double stamp);
int loam_lm_sync(loam_lm* h);
void* loam_lm_stream(loam_lm* h);
/* map persistence (the reference cannot save its map): checkpoint / resume of
 * the cube map laserCloudCornerArray / laserCloudSurfArray (laserMapping.cpp:74-81), the window
 * centre (:64-66), transformSum..transformAftMapped (:96-100), frame counters (:404-405) and the
 * IMU ring (:102-108) as one "LOAMMAP1" file (layout: csrc/lm_persist.cuh).  A run resumed from
 * the file continues bit-identically. */
int loam_lm_save_map(loam_lm* h, const char* path);
int loam_lm_load_map(loam_lm* h, const char* path);
/* a cloud of the node as binary PCD v0.7 (FIELDS x y z intensity): which = a loam_cloud selector
 * accepted by loam_lm_download, or 20 / 21 = the whole corner / surf cube map */
int loam_lm_export_pcd(loam_lm* h, int which, const char* path);
int loam_lm_download(loam_lm* h, int which_cloud, float* xyzi, int cap_points, int* n);
/* which = 100 / 101: the five neighbour indices per corner / surf query of the last LM iteration (-1 = rejected) */
int loam_lm_download_ints(loam_lm* h, int which, int* out, int cap, int* n);
int loam_lm_state(loam_lm* h, float aft_mapped[6], float bef_mapped[6], float tobe_mapped[6], int stats[8]);

/* ---- transformMaintenance node (transformMaintenance.cpp), host only ---- */
typedef struct loam_tm loam_tm;
int loam_tm_create(loam_tm** out);
void loam_tm_destroy(loam_tm* h);
/* odomAftMappedHandler :182-203 */
int loam_tm_on_mapping(loam_tm* h, const float aft_mapped[6], const float bef_mapped[6]);
/* laserOdometryHandler :147-180; writes /integrated_to_init */
int loam_tm_on_odometry(loam_tm* h, const float transform_sum[6], float integrated[6]);

/* the same two handlers for poses that were already decoded from their nav_msgs/Odometry messages
 * (transformMaintenance.cpp:151-161 / :186-196 in a ROS node) */
int loam_tm_on_mapping_msg(loam_tm* h, const float aft_mapped_decoded[6], const float bef_mapped[6]);
int loam_tm_on_odometry_msg(loam_tm* h, const float transform_sum_decoded[6], float integrated[6]);

/* ---- whole pipeline: sr -> lo -> (lm when lo publishes) -> tm, one stream, one device ---- */
typedef struct loam_pipeline loam_pipeline;
int loam_pipeline_create(const loam_params* p, int device, loam_pipeline** out);
void loam_pipeline_destroy(loam_pipeline* h);
int loam_pipeline_imu(loam_pipeline* h, double stamp, const double quat[4], const float acc[3]);
/* synchronous: uploads the sweep, runs the nodes, downloads the poses into *res */
int loam_pipeline_push(loam_pipeline* h, const float* xyz, int n, int stride_bytes, double stamp,
                       loam_sweep_result* res);
/* asynchronous form: push_async only enqueues.  A pageable (or strided) host buffer is staged
 * into pinned memory before it returns; a CUDA-pinned packed buffer (cudaHostAlloc /
 * cudaHostRegister, stride_bytes == 12) is read in place by the copy engine and must stay
 * unchanged until collect().  The H2D transfer runs on its own stream and overlaps the kernels
 * of the previous sweep.  collect() waits for everything enqueued so far and returns the per-sweep results in push order.
 * push_device takes a sweep that is already resident in device memory as packed xyz floats;
 * first_last_xyz = host copies of its first and last finite point (scanRegistration.cpp:230-232
 * reads them on the host), first_index = index of the first finite point. */
int loam_pipeline_push_async(loam_pipeline* h, const float* xyz, int n, int stride_bytes, double stamp);
int loam_pipeline_push_device(loam_pipeline* h, const float* d_xyz, int n, const float first_last_xyz[6],
                              int first_index, double stamp);
int loam_pipeline_collect(loam_pipeline* h, loam_sweep_result* out, int cap, int* n);
/* the synchronous and the enqueue-only push for a sweep that arrives as raw VLP-16 packets */
int loam_pipeline_push_packets(loam_pipeline* h, const unsigned char* packets, int n_packets, double stamp,
                               loam_sweep_result* res);
int loam_pipeline_push_packets_async(loam_pipeline* h, const unsigned char* packets, int n_packets, double stamp);
/* The stages of consecutive sweeps overlap on internal streams (registration of sweep k+1 and
 * mapping of sweep k-1 beside the odometry of sweep k).  loam_pipeline_stream() is the odometry
 * stream; loam_pipeline_join() makes it wait (on the device, no host synchronisation) for all work
 * enqueued so far on the other streams, so that an event recorded on it afterwards covers
 * everything pushed before. */
void* loam_pipeline_stream(loam_pipeline* h);
int loam_pipeline_join(loam_pipeline* h);
loam_sr* loam_pipeline_sr(loam_pipeline* h);
loam_lo* loam_pipeline_lo(loam_pipeline* h);
loam_lm* loam_pipeline_lm(loam_pipeline* h);

/* ---- many sequences on one GPU (BASELINE.json configs[4]) ----
 * `count` independent pipelines whose scanRegistration passes are issued as ONE launch chain per step (a batched
 * form of every kernel of scanRegistration.cpp:211-636, blockIdx.z = sequence); laserOdometry and laserMapping
 * stay per sequence.  push_async takes one sweep per sequence (xyz[i]: n[i] points, HOST memory, stride_bytes
 * apart) and only enqueues; results are collected per sequence with loam_pipeline_collect on
 * loam_pipeline_batch_get(h, i) (do not destroy those handles: the batch owns them).  Sequences with IMU data
 * are not supported here (use one loam_pipeline per sequence).  Not thread-safe. */
typedef struct loam_pipeline_batch loam_pipeline_batch;
int loam_pipeline_batch_create(const loam_params* p, int device, int count, loam_pipeline_batch** out);
void loam_pipeline_batch_destroy(loam_pipeline_batch* h);
int loam_pipeline_batch_size(loam_pipeline_batch* h);
loam_pipeline* loam_pipeline_batch_get(loam_pipeline_batch* h, int i);
int loam_pipeline_batch_push_async(loam_pipeline_batch* h, const float* const* xyz, const int* n, int stride_bytes,
                                   const double* stamps);
/* the same for sweeps that are already resident in device memory (the form of loam_pipeline_push_device) */
int loam_pipeline_batch_push_device(loam_pipeline_batch* h, const float* const* d_xyz, const int* n,
                                    const float* const* first_last_xyz, const int* first_index, const double* stamps);

/* ---- allocation-free entry point ---- */
/* allocation-free variant of the whole-cloud voxel filter.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same kernel and same bits as
 * loam_debug_voxel_filter_coop (pcl::VoxelGrid semantics), but the inputs and outputs
 * are DEVICE pointers owned by the caller (d_xyzi: n points x,y,z,intensity; d_out: room for
 * cap_out >= n points; *d_n_out: number of output points, a device int) and all scratch lives in one
 * static __device__ array inside the library (about 4.3 MB, held by every GPU that loads the library
 * until exit), so the call performs no cudaMalloc / cudaFree / allocating library routine.  It only
 * enqueues on `stream` (a cudaStream_t, NULL = the default stream) of the CURRENT device.
 * n > LOAM_NOALLOC_MAX_POINTS returns LOAM_ERR_CAPACITY with a message in loam_last_error().
 * All calls share the scratch array: calls that overlap on the GPU (different streams or host
 * threads) race -- serialise them. */
#define LOAM_NOALLOC_MAX_POINTS 131072
int loam_voxel_filter_noalloc(const float* d_xyzi, int n, float leaf, float* d_out, int cap_out, int* d_n_out,
                              void* stream);

/* test / development hooks (loam_debug_*): include/loam_b200_debug.h */

/* number of kernel launches issued by this library on the calling process so far */
long long loam_kernel_launch_count(void);
/* per-kernel device time accumulated since the last reset (CUDA events; 0 when profiling is off) */
int loam_profile_enable(int on);
int loam_profile_report(char* buf, int cap);
const char* loam_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* LOAM_B200_H_ */
