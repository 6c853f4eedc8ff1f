// ros_mirror.hpp — ROS-free stand-ins for the message types PlanningState touches
// (sensor_msgs/JointState, trajectory_msgs/JointTrajectory[Point], occupancy_grid/OccupancyGrid,
// ros::Publisher), field for field as far as the path uses them.  With -DDOAPP_WITH_ROS the real
// headers are used instead and these names alias the ROS types.
#pragma once

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#ifdef DOAPP_WITH_ROS
#include "ros/ros.h"
#include "sensor_msgs/JointState.h"
#include "trajectory_msgs/JointTrajectory.h"
namespace doapp_host {
using JointState = sensor_msgs::JointState;
using JointTrajectoryPoint = trajectory_msgs::JointTrajectoryPoint;
using JointTrajectory = trajectory_msgs::JointTrajectory;
using Publisher = ros::Publisher;
inline double to_sec(const ros::Duration &d) { return d.toSec(); }
inline ros::Duration from_sec(double s) { return ros::Duration(s); }
} // namespace doapp_host
#else
namespace doapp_host {

struct Header {
  uint32_t seq = 0;
  double stamp = 0.0; // seconds
  std::string frame_id;
};

// sensor_msgs/JointState: only .position is read (planning_state.cuh:30-44)
struct JointState {
  using ConstPtr = std::shared_ptr<const JointState>;
  Header header;
  std::vector<std::string> name;
  std::vector<double> position, velocity, effort;
};

// trajectory_msgs/JointTrajectoryPoint: positions, velocities, time_from_start are written
// (planning_state.cu:179-209)
struct JointTrajectoryPoint {
  std::vector<double> positions, velocities, accelerations, effort;
  double time_from_start = 0.0; // seconds
};

struct JointTrajectory {
  Header header;
  std::vector<std::string> joint_names; // left empty by the reference
  std::vector<JointTrajectoryPoint> points;
};

// ros::Publisher stand-in: keeps what was published so a test can read it back
struct Publisher {
  std::vector<JointTrajectory> published;
  void publish(JointTrajectory msg) { published.push_back(std::move(msg)); }
};

inline double to_sec(double d) { return d; }
inline double from_sec(double s) { return s; }

} // namespace doapp_host
#endif

namespace doapp_host {
// occupancy_grid/msg/OccupancyGrid.msg:1-3
struct OccupancyGridMsg {
  double size = 0.0;
  double granularity = 0.0;
  std::vector<int8_t> grid;
};
// the xyz payload of a /merged_cloud-shaped message (mapping_node.cpp:241-256,395)
struct PointCloudXYZ {
  std::vector<float> xyz; // [P][3]
};
} // namespace doapp_host
