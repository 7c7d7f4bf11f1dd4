// scan_registration_node.cpp -- ROS shell of the reference's scanRegistration node
// (src/scanRegistration.cpp:662-700: same node name, same two subscriptions, same six
// publications) around libloam_b200: the numeric body of laserCloudHandler (:211-636) and of
// imuHandler (:638-660) is one C-ABI call each.
#include "adapter_common.h"

using namespace loam_ros;

static loam_sr* g_sr = 0;
static int g_system_delay = 20;   // scanRegistration.cpp:57
static ros::Publisher g_pub[5], g_pub_imu_trans;
static std::vector<float> g_xyz, g_out;

static void laserCloudHandler(const sensor_msgs::PointCloud2ConstPtr& msg) {
  pcl::PointCloud<pcl::PointXYZ> in;
  pcl::fromROSMsg(*msg, in);
  const size_t n = in.points.size();
  g_xyz.resize(3 * n + 3);
  for (size_t i = 0; i < n; i++) { g_xyz[3 * i] = in.points[i].x; g_xyz[3 * i + 1] = in.points[i].y; g_xyz[3 * i + 2] = in.points[i].z; }
  const double stamp = msg->header.stamp.toSec();
  int rc = loam_sr_process(g_sr, g_xyz.data(), (int)n, 12, stamp);
  die_on_error(rc, "loam_sr_process");
  if (rc != LOAM_OK) return;   // the sweeps the reference swallows (:213-219) publish nothing
  static const int which[5] = {LOAM_CLOUD_FULL, LOAM_CLOUD_SHARP, LOAM_CLOUD_LESS_SHARP, LOAM_CLOUD_FLAT, LOAM_CLOUD_LESS_FLAT};
  g_out.resize(4 * n + 4);
  for (int c = 0; c < 5; c++) {
    int cnt = 0;
    die_on_error(loam_sr_download(g_sr, which[c], g_out.data(), (int)n, &cnt), "loam_sr_download");
    publish_cloud(g_pub[c], g_out, cnt, stamp, "/camera");
  }
  float it[12];
  die_on_error(loam_sr_imu_trans(g_sr, it), "loam_sr_imu_trans");
  pcl::PointCloud<pcl::PointXYZ> imuTrans;   // four xyz points, :614-629
  imuTrans.points.resize(4);
  for (int i = 0; i < 4; i++) { imuTrans.points[i].x = it[3 * i]; imuTrans.points[i].y = it[3 * i + 1]; imuTrans.points[i].z = it[3 * i + 2]; }
  imuTrans.width = 4; imuTrans.height = 1;
  sensor_msgs::PointCloud2 m;
  pcl::toROSMsg(imuTrans, m);
  m.header.stamp = msg->header.stamp;
  m.header.frame_id = "/camera";
  g_pub_imu_trans.publish(m);
}

static void imuHandler(const sensor_msgs::Imu::ConstPtr& imu) {
  const double q[4] = {imu->orientation.x, imu->orientation.y, imu->orientation.z, imu->orientation.w};
  const float acc[3] = {(float)imu->linear_acceleration.x, (float)imu->linear_acceleration.y, (float)imu->linear_acceleration.z};
  die_on_error(loam_sr_imu(g_sr, imu->header.stamp.toSec(), q, acc), "loam_sr_imu");
}

int main(int argc, char** argv) {
  ros::init(argc, argv, "scanRegistration");
  ros::NodeHandle nh;
  loam_params p = params_from_env(g_system_delay);
  die_on_error(loam_sr_create(&p, env_int("LOAM_B200_DEVICE", 0), &g_sr), "loam_sr_create");
  ros::Subscriber subCloud = nh.subscribe<sensor_msgs::PointCloud2>("/velodyne_points", 2, laserCloudHandler);
  ros::Subscriber subImu = nh.subscribe<sensor_msgs::Imu>("/imu/data", 50, imuHandler);
  static const char* topics[5] = {"/velodyne_cloud_2", "/laser_cloud_sharp", "/laser_cloud_less_sharp", "/laser_cloud_flat", "/laser_cloud_less_flat"};
  for (int c = 0; c < 5; c++) g_pub[c] = nh.advertise<sensor_msgs::PointCloud2>(topics[c], 2);
  g_pub_imu_trans = nh.advertise<sensor_msgs::PointCloud2>("/imu_trans", 5);
  ros::spin();
  loam_sr_destroy(g_sr);
  return 0;
}
