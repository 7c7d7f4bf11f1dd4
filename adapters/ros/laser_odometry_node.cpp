// laser_odometry_node.cpp -- ROS shell of the reference's laserOdometry node
// (src/laserOdometry.cpp:353-931: six subscriptions, 5 ms stamp synchronisation, four
// publications + tf) around libloam_b200: the loop body (:427-911) is loam_lo_process_host.
#include "adapter_common.h"

using namespace loam_ros;

static loam_lo* g_lo = 0;
static std::vector<float> g_cloud[5];          // full, sharp, less sharp, flat, less flat
static double g_time[6] = {0, 0, 0, 0, 0, 0};  // ... and imu_trans
static bool g_new[6] = {false, false, false, false, false, false};
static float g_imu_trans[12];
static float g_transform[6], g_transform_sum[6];

template <int K> static void cloudHandler(const sensor_msgs::PointCloud2ConstPtr& msg) {
  g_time[K] = msg->header.stamp.toSec();
  unpack_msg(*msg, &g_cloud[K], true);   // removeNaNFromPointCloud, laserOdometry.cpp:282-326
  g_new[K] = true;
}
static void imuTransHandler(const sensor_msgs::PointCloud2ConstPtr& msg) {   // :328-351
  g_time[5] = msg->header.stamp.toSec();
  pcl::PointCloud<pcl::PointXYZ> t;
  pcl::fromROSMsg(*msg, t);
  for (int i = 0; i < 4 && i < (int)t.points.size(); i++) { g_imu_trans[3 * i] = t.points[i].x; g_imu_trans[3 * i + 1] = t.points[i].y; g_imu_trans[3 * i + 2] = t.points[i].z; }
  g_new[5] = true;
}

int main(int argc, char** argv) {
  ros::init(argc, argv, "laserOdometry");
  ros::NodeHandle nh;
  loam_params p = params_from_env(0);
  die_on_error(loam_lo_create(&p, env_int("LOAM_B200_DEVICE", 0), &g_lo), "loam_lo_create");
  ros::Subscriber s1 = nh.subscribe<sensor_msgs::PointCloud2>("/laser_cloud_sharp", 2, cloudHandler<1>);
  ros::Subscriber s2 = nh.subscribe<sensor_msgs::PointCloud2>("/laser_cloud_less_sharp", 2, cloudHandler<2>);
  ros::Subscriber s3 = nh.subscribe<sensor_msgs::PointCloud2>("/laser_cloud_flat", 2, cloudHandler<3>);
  ros::Subscriber s4 = nh.subscribe<sensor_msgs::PointCloud2>("/laser_cloud_less_flat", 2, cloudHandler<4>);
  ros::Subscriber s0 = nh.subscribe<sensor_msgs::PointCloud2>("/velodyne_cloud_2", 2, cloudHandler<0>);
  ros::Subscriber s5 = nh.subscribe<sensor_msgs::PointCloud2>("/imu_trans", 5, imuTransHandler);
  ros::Publisher pubCorner = nh.advertise<sensor_msgs::PointCloud2>("/laser_cloud_corner_last", 2);
  ros::Publisher pubSurf = nh.advertise<sensor_msgs::PointCloud2>("/laser_cloud_surf_last", 2);
  ros::Publisher pubFull = nh.advertise<sensor_msgs::PointCloud2>("/velodyne_cloud_3", 2);
  ros::Publisher pubOdom = nh.advertise<nav_msgs::Odometry>("/laser_odom_to_init", 5);
  nav_msgs::Odometry odom;
  odom.header.frame_id = "/camera_init";
  odom.child_frame_id = "/laser_odom";
  tf::TransformBroadcaster br;
  tf::StampedTransform tr;
  tr.frame_id_ = "/camera_init";
  tr.child_frame_id_ = "/laser_odom";
  std::vector<float> out;
  ros::Rate rate(100);
  while (ros::ok()) {
    ros::spinOnce();
    bool ready = true;
    for (int k = 0; k < 6; k++) ready = ready && g_new[k] && fabs(g_time[k] - g_time[4]) < 0.005;   // :413-419
    if (ready) {
      for (int k = 0; k < 6; k++) g_new[k] = false;
      const double stamp = g_time[4];
      int pub = 0;
      int rc = loam_lo_process_host(g_lo, g_cloud[0].data(), (int)g_cloud[0].size() / 4, g_cloud[1].data(), (int)g_cloud[1].size() / 4,
                                    g_cloud[2].data(), (int)g_cloud[2].size() / 4, g_cloud[3].data(), (int)g_cloud[3].size() / 4,
                                    g_cloud[4].data(), (int)g_cloud[4].size() / 4, g_imu_trans, &pub);
      die_on_error(rc, "loam_lo_process_host");
      int st[4];
      die_on_error(loam_lo_state(g_lo, g_transform, g_transform_sum, st), "loam_lo_state");
      const bool first = rc == LOAM_NOT_READY;   // :427-456: the first sweep only hands its clouds on
      if (!first) {
        odom.header.stamp = ros::Time().fromSec(stamp);
        pose_to_msg(g_transform_sum, &odom);
        pubOdom.publish(odom);
        send_tf(br, tr, odom, stamp);
      }
      if (first || pub) {
        size_t cap = g_cloud[0].size() / 4 + g_cloud[2].size() / 4 + g_cloud[4].size() / 4 + 16;
        out.resize(4 * cap);
        int n = 0;
        die_on_error(loam_lo_download(g_lo, LOAM_CLOUD_CORNER_LAST, out.data(), (int)cap, &n), "loam_lo_download");
        publish_cloud(pubCorner, out, n, stamp, "/camera");
        die_on_error(loam_lo_download(g_lo, LOAM_CLOUD_SURF_LAST, out.data(), (int)cap, &n), "loam_lo_download");
        publish_cloud(pubSurf, out, n, stamp, "/camera");
        if (pub) {
          die_on_error(loam_lo_download(g_lo, LOAM_CLOUD_FULL_END, out.data(), (int)cap, &n), "loam_lo_download");
          publish_cloud(pubFull, out, n, stamp, "/camera");
        }
      }
    }
    rate.sleep();
  }
  loam_lo_destroy(g_lo);
  return 0;
}
