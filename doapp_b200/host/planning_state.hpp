// planning_state.hpp — the reference's PlanningState (perturbator/include/planning_state.cuh:17-73)
// re-hosted on the C ABI of include/doapp/doapp.h.  Same constructor arguments, same method
// names and meaning, same error behaviour (std::runtime_error from the constructor like
// vector.cuh:149-151 / cpu_distance_grid.cu:48-51; failures after construction abort through
// DOAPP_ASSERT like ROS_ASSERT(false), planning_state.cu:37-43).  New, beyond the reference:
// updateOccupancyGrid / updatePointCloud (the grid feed the reference never wired)
// and setIterations (the reference optimises until a host flag flips, gt.cu:253).
#pragma once

#include <cstddef>
#include <cstdint>
#include <vector>

#include "doapp/doapp.h"
#include "ros_mirror.hpp"

namespace doapp {
namespace cpu_distance_grid {
// perturbator/include/cpu_distance_grid.cuh:13-18
struct Dimensions {
  std::size_t length; // x
  std::size_t width;  // y
  std::size_t height; // z
  double resolution;  // meters
};
} // namespace cpu_distance_grid
} // namespace doapp

class PlanningState {
public:
  // planning_state.cu:9-44.  k climbers x m noise vectors x n waypoints x d joints.
  PlanningState(unsigned int num_initial_trajectories_, unsigned int num_noise_vectors_,
                unsigned int num_waypoints_, unsigned int waypoint_dim_,
                const doapp::cpu_distance_grid::Dimensions &dims);
  ~PlanningState();
  PlanningState(const PlanningState &) = delete;
  PlanningState &operator=(const PlanningState &) = delete;

  // planning_state.cuh:30-34: positions truncated / zero-padded to waypoint_dim
  void updateArmState(const doapp_host::JointState::ConstPtr &msg);
  // planning_state.cuh:37-44: no-op if the goal did not change
  void updateGoalState(const doapp_host::JointState::ConstPtr &msg);
  bool have_goal() { return _have_goal; }
  // planning_state.cu:56-90: asynchronous; returns after enqueueing the optimiser
  void plan();
  // planning_state.cu:108-212: waits, picks the best trajectory, publishes n + 2 points
  void publish_trajectory(doapp_host::Publisher &pub);
  // planning_state.cu:46-54 (with the missing write restored)
  std::vector<float> plan_lerp() const;

  // dense host distance grid (fp32, index ix + L*iy + L*W*iz) -> device; replaces the
  // KD-tree fill of updateDistanceGrid (planning_state.cuh:25-28, cpu_distance_grid.cu:87-202)
  void updateDistanceGrid(const std::vector<float> &distances);
  // the grid feed the reference never wired: an OccupancyGrid message, or an xyz cloud
  // scattered into a uint8 occupancy grid of the constructor's dimensions
  void updateOccupancyGrid(const doapp_host::OccupancyGridMsg &msg);
  void updatePointCloud(const doapp_host::PointCloudXYZ &cloud, bool clear_first = true);
  // depth frames -> occupancy in one kernel (mapping_node.cpp:89-137,195,262-263): see
  // doapp_grid_from_depth for the argument layout
  void updateDepthFrames(const uint16_t *depth, unsigned int num_cams, unsigned int width,
                         unsigned int height, const double *intrinsics, const float *extrinsics,
                         double crop_size, bool clear_first = true);
  // occupancy -> exact Euclidean distance field: what updateDistanceGrid(KDTree) was meant to
  // produce on the host (cpu_distance_grid.cu:87-202); the sphere test of gt.cu:412 then applies
  void buildDistanceField();
  // collision term on the arm's real kinematic chain (rexarm.urdf:61-149, 13 spheres along the
  // link boxes) instead of the point chain of generate_trajectories.cu:397-463
  void useUrdfKinematics(bool on);
  // The follower indexes positions[i] / velocities[i] for i < 6 servo ids
  // (dynamixel_driver/src/trajectory_follower.cpp:176-178,194-199) while the planner emits d
  // entries per point: with n > d the published points are padded to n entries,
  // holding the extra servos at their last reported position with zero velocity.
  void setFollowerJoints(unsigned int n) { follower_joints = n; }
  void setIterations(unsigned int iters) { iterations = iters; }
  float last_cost() const { return last_best_score; }
  doapp_ctx *handle() { return ctx; }

private:
  void synchronize();
  const float deltaT = 0.1f; // planning_state.cuh:55
  bool planning_trajectory = false, _have_goal = false;
  unsigned int num_initial_trajectories, num_noise_vectors, num_waypoints, waypoint_dim;
  unsigned int iterations = 100;
  std::vector<float> goal_state, current_state;
  doapp::cpu_distance_grid::Dimensions dims;
  doapp_ctx *ctx = nullptr;
  float last_best_score = 0.0f;
  bool have_u8_grid = false;
  unsigned int follower_joints = 0;
  std::vector<float> reported_state; // every joint of the last joint_state message
};
