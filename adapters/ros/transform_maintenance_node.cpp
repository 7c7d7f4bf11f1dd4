// transform_maintenance_node.cpp -- ROS shell of the reference's transformMaintenance node
// (src/transformMaintenance.cpp:147-235): host only, as in the reference.
#include "adapter_common.h"

using namespace loam_ros;

static loam_tm* g_tm = 0;
static ros::Publisher* g_pub = 0;
static tf::TransformBroadcaster* g_br = 0;
static nav_msgs::Odometry g_out;
static tf::StampedTransform g_tr;
static float g_mapped[6];

static void laserOdometryHandler(const nav_msgs::Odometry::ConstPtr& msg) {   // :147-180
  float sum[6];
  pose_from_msg(*msg, sum);
  die_on_error(loam_tm_on_odometry_msg(g_tm, sum, g_mapped), "loam_tm_on_odometry_msg");
  g_out.header.stamp = msg->header.stamp;
  pose_to_msg(g_mapped, &g_out);
  g_pub->publish(g_out);
  send_tf(*g_br, g_tr, g_out, msg->header.stamp.toSec());
}
static void odomAftMappedHandler(const nav_msgs::Odometry::ConstPtr& msg) {   // :182-203
  float aft[6], bef[6];
  pose_from_msg(*msg, aft);
  bef[0] = msg->twist.twist.angular.x; bef[1] = msg->twist.twist.angular.y; bef[2] = msg->twist.twist.angular.z;
  bef[3] = msg->twist.twist.linear.x; bef[4] = msg->twist.twist.linear.y; bef[5] = msg->twist.twist.linear.z;
  die_on_error(loam_tm_on_mapping_msg(g_tm, aft, bef), "loam_tm_on_mapping_msg");
}

int main(int argc, char** argv) {
  ros::init(argc, argv, "transformMaintenance");
  ros::NodeHandle nh;
  die_on_error(loam_tm_create(&g_tm), "loam_tm_create");
  ros::Subscriber s0 = nh.subscribe<nav_msgs::Odometry>("/laser_odom_to_init", 5, laserOdometryHandler);
  ros::Subscriber s1 = nh.subscribe<nav_msgs::Odometry>("/aft_mapped_to_init", 5, odomAftMappedHandler);
  ros::Publisher pub = nh.advertise<nav_msgs::Odometry>("/integrated_to_init", 5);
  g_pub = &pub;
  g_out.header.frame_id = "/camera_init";
  g_out.child_frame_id = "/camera";
  tf::TransformBroadcaster br;
  g_br = &br;
  g_tr.frame_id_ = "/camera_init";
  g_tr.child_frame_id_ = "/camera";
  ros::spin();
  loam_tm_destroy(g_tm);
  return 0;
}
