// laser_mapping_node.cpp -- ROS shell of the reference's laserMapping node
// (src/laserMapping.cpp:337-1096: five subscriptions, 5 ms stamp synchronisation, three
// publications + tf) around libloam_b200: the loop body (:420-1069) is loam_lm_process_msg.
#include "adapter_common.h"

using namespace loam_ros;

static loam_lm* g_lm = 0;
static std::vector<float> g_cloud[3];   // corner last, surf last, full resolution
static double g_time[4] = {0, 0, 0, 0}; // ... and the odometry
static bool g_new[4] = {false, false, false, false};
static float g_transform_sum[6];
static float g_aft[6], g_bef[6], g_tobe[6];
static int g_stats[8];

template <int K> static void cloudHandler(const sensor_msgs::PointCloud2ConstPtr& msg) {   // :274-301 (no NaN filter here)
  g_time[K] = msg->header.stamp.toSec();
  unpack_msg(*msg, &g_cloud[K], false);
  g_new[K] = true;
}
static void laserOdometryHandler(const nav_msgs::Odometry::ConstPtr& msg) {   // :303-321
  g_time[3] = msg->header.stamp.toSec();
  pose_from_msg(*msg, g_transform_sum);
  g_new[3] = true;
}
static void imuHandler(const sensor_msgs::Imu::ConstPtr& imu) {   // :323-335
  const double q[4] = {imu->orientation.x, imu->orientation.y, imu->orientation.z, imu->orientation.w};
  die_on_error(loam_lm_imu(g_lm, imu->header.stamp.toSec(), q), "loam_lm_imu");
}

int main(int argc, char** argv) {
  ros::init(argc, argv, "laserMapping");
  ros::NodeHandle nh;
  loam_params p = params_from_env(0);
  die_on_error(loam_lm_create(&p, env_int("LOAM_B200_DEVICE", 0), &g_lm), "loam_lm_create");
  ros::Subscriber s0 = nh.subscribe<sensor_msgs::PointCloud2>("/laser_cloud_corner_last", 2, cloudHandler<0>);
  ros::Subscriber s1 = nh.subscribe<sensor_msgs::PointCloud2>("/laser_cloud_surf_last", 2, cloudHandler<1>);
  ros::Subscriber s3 = nh.subscribe<nav_msgs::Odometry>("/laser_odom_to_init", 5, laserOdometryHandler);
  ros::Subscriber s2 = nh.subscribe<sensor_msgs::PointCloud2>("/velodyne_cloud_3", 2, cloudHandler<2>);
  ros::Subscriber s4 = nh.subscribe<sensor_msgs::Imu>("/imu/data", 50, imuHandler);
  ros::Publisher pubSurround = nh.advertise<sensor_msgs::PointCloud2>("/laser_cloud_surround", 1);
  ros::Publisher pubRegistered = nh.advertise<sensor_msgs::PointCloud2>("/velodyne_cloud_registered", 2);
  ros::Publisher pubAftMapped = nh.advertise<nav_msgs::Odometry>("/aft_mapped_to_init", 5);
  nav_msgs::Odometry aft;
  aft.header.frame_id = "/camera_init";
  aft.child_frame_id = "/aft_mapped";
  tf::TransformBroadcaster br;
  tf::StampedTransform tr;
  tr.frame_id_ = "/camera_init";
  tr.child_frame_id_ = "/aft_mapped";
  std::vector<float> out;
  int passes = 0;
  ros::Rate rate(100);
  while (ros::ok()) {
    ros::spinOnce();
    bool ready = true;
    for (int k = 0; k < 4; k++) ready = ready && g_new[k];
    ready = ready && fabs(g_time[0] - g_time[3]) < 0.005 && fabs(g_time[1] - g_time[3]) < 0.005 && fabs(g_time[2] - g_time[3]) < 0.005;   // :409-414
    if (ready) {
      for (int k = 0; k < 4; k++) g_new[k] = false;
      const double stamp = g_time[3];
      int rc = loam_lm_process_msg(g_lm, g_cloud[0].data(), (int)g_cloud[0].size() / 4, g_cloud[1].data(), (int)g_cloud[1].size() / 4,
                                   g_cloud[2].data(), (int)g_cloud[2].size() / 4, g_transform_sum, stamp);
      die_on_error(rc, "loam_lm_process_msg");
      if (rc == LOAM_OK) {
        die_on_error(loam_lm_state(g_lm, g_aft, g_bef, g_tobe, g_stats), "loam_lm_state");
        passes++;
        int n = 0;
        if (passes % 5 == 1) {   // mapFrameNum = 5, counter primed to publish on the first pass (:52, :405, :1038-1058)
          die_on_error(loam_lm_download(g_lm, LOAM_CLOUD_SURROUND, 0, 0, &n), "loam_lm_download");
          out.resize(4 * (size_t)n + 4);
          die_on_error(loam_lm_download(g_lm, LOAM_CLOUD_SURROUND, out.data(), n, &n), "loam_lm_download");
          publish_cloud(pubSurround, out, n, stamp, "/camera_init");
        }
        out.resize(g_cloud[2].size() + 4);
        die_on_error(loam_lm_download(g_lm, LOAM_CLOUD_REGISTERED, out.data(), (int)g_cloud[2].size() / 4, &n), "loam_lm_download");
        publish_cloud(pubRegistered, out, n, stamp, "/camera_init");
        aft.header.stamp = ros::Time().fromSec(stamp);
        pose_to_msg(g_aft, &aft);
        aft.twist.twist.angular.x = g_bef[0]; aft.twist.twist.angular.y = g_bef[1]; aft.twist.twist.angular.z = g_bef[2];   // :1082-1087
        aft.twist.twist.linear.x = g_bef[3]; aft.twist.twist.linear.y = g_bef[4]; aft.twist.twist.linear.z = g_bef[5];
        pubAftMapped.publish(aft);
        send_tf(br, tr, aft, stamp);
      }
    }
    rate.sleep();
  }
  loam_lm_destroy(g_lm);
  return 0;
}
