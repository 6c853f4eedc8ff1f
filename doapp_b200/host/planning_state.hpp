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

// The reference's updateDistanceGrid takes a nanoflann KD-tree over the obstacle points
// (planning_state.cuh:25-28, cpu_distance_grid.cuh:50-52).  nanoflann is a header-only third
// party library vendored by the reference (perturbator/nanoflann, 1.3.2): when it is on the
// include path the same overload exists here.
#if !defined(DOAPP_WITH_NANOFLANN) && defined(__has_include)
#if __has_include(<nanoflann.hpp>)
#define DOAPP_WITH_NANOFLANN 1
#endif
#endif
#ifdef DOAPP_WITH_NANOFLANN
#include <nanoflann.hpp>
#endif

namespace doapp {
namespace cpu_distance_grid {
// perturbator/include/cpu_distance_grid.cuh:13-18
struct Dimensions {
  std::size_t length; // x
  std::size_t width;  // y
  std::size_t height; // z
  double resolution;  // meters
};
#ifdef DOAPP_WITH_NANOFLANN
// Dataset adaptor with the interface of cpu_distance_grid.cuh:20-48, over a plain [P][3] array
// (the reference's adaptor wraps its own managed-memory Matrix<float, Dynamic, 3>).
class MatrixPointcloudAdaptor {
public:
  using coord_t = float;
  explicit MatrixPointcloudAdaptor(const std::vector<float> &xyz) noexcept : xyz_(&xyz) {}
  std::size_t kdtree_get_point_count() const noexcept { return xyz_->size() / 3; }
  float kdtree_get_pt(std::size_t index, std::size_t dimension) const { return (*xyz_)[3 * index + dimension]; }
  template <typename BoundingBox> bool kdtree_get_bbox(BoundingBox &) const noexcept { return false; }

private:
  const std::vector<float> *xyz_;
};
// cpu_distance_grid.cuh:50-52
using KDTree = nanoflann::KDTreeSingleIndexAdaptor<
    nanoflann::L2_Simple_Adaptor<float, MatrixPointcloudAdaptor>, MatrixPointcloudAdaptor, 3>;
#endif
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

#ifdef DOAPP_WITH_NANOFLANN
  // planning_state.cuh:25-28, same signature.  The reference fills the distance grid on the host
  // with one nearest-neighbour query per voxel (cpu_distance_grid.cu:87-202, 10^6 KD-tree
  // queries for its 100^3 grid); here the tree's points are scattered into the occupancy grid
  // and the exact Euclidean distance transform runs on the device (doapp_grid_scatter +
  // doapp_grid_distance_field) -- distances between voxel centres instead of voxel centre to
  // raw point, at most half a voxel diagonal apart.
  void updateDistanceGrid(const doapp::cpu_distance_grid::KDTree &tree) noexcept {
    const auto &pts = tree.dataset;
    doapp_host::PointCloudXYZ cloud;
    const std::size_t count = pts.kdtree_get_point_count();
    cloud.xyz.resize(3 * count);
    for (std::size_t i = 0; i < count; ++i)
      for (std::size_t a = 0; a < 3; ++a) cloud.xyz[3 * i + a] = pts.kdtree_get_pt(i, a);
    updatePointCloud(cloud, true);
    buildDistanceField();
  }
#endif
  // dense host distance grid (fp32, index ix + L*iy + L*W*iz) -> device (a caller that already
  // holds the distances; not in the reference)
  void uploadDistanceGrid(const std::vector<float> &distances);
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
