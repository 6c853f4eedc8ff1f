// planning_state.cpp — PlanningState over the C ABI (see planning_state.hpp).
#include "planning_state.hpp"

#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>

// ROS_ASSERT(false) stand-in: the reference aborts on any CUDA error after construction
// (planning_state.cu:37-43,63-77,85-88,96-105)
#ifdef DOAPP_WITH_ROS
#define DOAPP_ASSERT(cond) ROS_ASSERT(cond)
#else
#define DOAPP_ASSERT(cond)                                                              \
  do {                                                                                  \
    if (!(cond)) {                                                                      \
      std::fprintf(stderr, "ASSERTION FAILED %s:%d: %s (%s)\n", __FILE__, __LINE__, #cond, \
                   ctx ? doapp_last_error(ctx) : "");                                   \
      std::abort();                                                                     \
    }                                                                                   \
  } while (0)
#endif

PlanningState::PlanningState(unsigned int num_initial_trajectories_,
                             unsigned int num_noise_vectors_, unsigned int num_waypoints_,
                             unsigned int waypoint_dim_,
                             const doapp::cpu_distance_grid::Dimensions &dims_)
    : num_initial_trajectories(num_initial_trajectories_), num_noise_vectors(num_noise_vectors_),
      num_waypoints(num_waypoints_), waypoint_dim(waypoint_dim_), goal_state(waypoint_dim_, 0.0f),
      current_state(waypoint_dim_, 0.0f), dims(dims_) {
  doapp_config cfg;
  doapp_config_init(&cfg);
  cfg.num_climbers = num_initial_trajectories;
  cfg.samples_per_climber = num_noise_vectors;
  cfg.num_waypoints = num_waypoints;
  cfg.waypoint_dim = waypoint_dim;
  cfg.delta_t = deltaT;
  cfg.select_mode = DOAPP_SELECT_CEIL_INT; // int(ceilf(score)) + atomicMin, gt.cu:109-111
  if (doapp_create(&cfg, &ctx) != DOAPP_OK) // vector.cuh:149-151 / cdg.cu:48-51 throw here too
    throw std::runtime_error(std::string("PlanningState::PlanningState: ") + doapp_last_error(nullptr));
  // CPUDistanceGrid ctor (cdg.cu:28-83): x,y centred, z from 0, "nothing nearby" everywhere
  if (doapp_grid_alloc(ctx, DOAPP_GRID_F32, DOAPP_FRAME_CDG, (uint32_t)dims.length,
                       (uint32_t)dims.width, (uint32_t)dims.height, dims.resolution) != DOAPP_OK) {
    std::string e = doapp_last_error(ctx);
    doapp_destroy(ctx);
    ctx = nullptr;
    throw std::runtime_error("PlanningState::PlanningState: " + e);
  }
}

PlanningState::~PlanningState() { doapp_destroy(ctx); }

void PlanningState::updateArmState(const doapp_host::JointState::ConstPtr &msg) {
  current_state = std::vector<float>(msg->position.begin(), msg->position.end());
  reported_state = current_state;
  current_state.resize(waypoint_dim); // planning_state.cuh:32 (num_joints there, d here: D3)
  DOAPP_ASSERT(current_state.size() == waypoint_dim);
}

void PlanningState::updateGoalState(const doapp_host::JointState::ConstPtr &msg) {
  std::vector<float> new_goal(msg->position.begin(), msg->position.end());
  new_goal.resize(waypoint_dim);
  if (new_goal == goal_state) return; // planning_state.cuh:39
  _have_goal = true;
  goal_state = new_goal;
}

std::vector<float> PlanningState::plan_lerp() const {
  std::vector<float> result(num_waypoints * waypoint_dim);
  doapp_plan_lerp(current_state.data(), goal_state.data(), num_waypoints, waypoint_dim,
                  result.data());
  return result;
}

void PlanningState::plan() {
  if (!_have_goal) return;
  synchronize(); // planning_state.cu:59: stop whatever is still running
  std::vector<float> lerp = plan_lerp();
  doapp_status st = doapp_plan_async(ctx, current_state.data(), goal_state.data(), lerp.data(),
                                     iterations, nullptr);
  DOAPP_ASSERT(st == DOAPP_OK);
  planning_trajectory = true;
}

void PlanningState::synchronize() {
  if (!planning_trajectory) return;
  // the reference flips a managed flag and cudaDeviceSynchronize()s (planning_state.cu:92-106);
  // here the launch is bounded, waiting for it is enough
  doapp_status st = doapp_plan_wait(ctx, nullptr, nullptr, nullptr, nullptr);
  DOAPP_ASSERT(st == DOAPP_OK || st == DOAPP_NOT_READY);
  planning_trajectory = false;
}

void PlanningState::publish_trajectory(doapp_host::Publisher &pub) {
  if (!planning_trajectory) return;
  const size_t n = num_waypoints, d = waypoint_dim;
  std::vector<float> traj(n * d), vel(n * d);
  long long climber = -1;
  doapp_status st = doapp_plan_wait(ctx, traj.data(), vel.data(), &last_best_score,
                                    reinterpret_cast<int64_t *>(&climber));
  DOAPP_ASSERT(st == DOAPP_OK);
  planning_trajectory = false;
  (void)climber; // -1: no climber improved; best_traj is then the initial (lerp) trajectory

  doapp_host::JointTrajectory msg; // planning_state.cu:179-209
#ifdef DOAPP_WITH_ROS
  msg.header.stamp = ros::Time::now();
#endif
  doapp_host::JointTrajectoryPoint point;
  for (size_t j = 0; j < d; ++j) { // current state, velocity towards waypoint 0
    point.positions.push_back(current_state[j]);
    point.velocities.push_back((traj[j] - current_state[j]) / deltaT);
  }
  point.time_from_start = doapp_host::from_sec(0);
  msg.points.push_back(point);
  for (size_t i = 0; i < n; ++i) {
    point.positions.clear();
    point.velocities.clear();
    for (size_t j = 0; j < d; ++j) {
      point.positions.push_back(traj[i * d + j]);
      // look-forward velocity (planning_state.cu:195-197); the read past the last waypoint
      // there is a zero here
      point.velocities.push_back(i + 1 < n ? vel[(i + 1) * d + j] : 0.0f);
    }
    point.time_from_start = doapp_host::from_sec(deltaT * (float)i);
    msg.points.push_back(point);
  }
  point.positions.clear();
  point.velocities.clear();
  for (size_t j = 0; j < d; ++j) { // goal, stop at the last point
    point.positions.push_back(goal_state[j]);
    point.velocities.push_back(0);
  }
  msg.points.push_back(point); // keeps the last waypoint's stamp, as in the reference
  if (follower_joints > d) // trajectory_follower.cpp:176-178 reads 6 entries per point
    for (auto &pt : msg.points)
      for (size_t j = d; j < follower_joints; ++j) {
        pt.positions.push_back(j < reported_state.size() ? reported_state[j] : 0.0f);
        pt.velocities.push_back(0);
      }
  pub.publish(std::move(msg));
  _have_goal = false;
}

void PlanningState::uploadDistanceGrid(const std::vector<float> &distances) {
  synchronize();
  doapp_status st = doapp_grid_alloc(ctx, DOAPP_GRID_F32, DOAPP_FRAME_CDG, (uint32_t)dims.length,
                                     (uint32_t)dims.width, (uint32_t)dims.height, dims.resolution);
  DOAPP_ASSERT(st == DOAPP_OK);
  st = doapp_grid_upload(ctx, distances.data(), distances.size());
  DOAPP_ASSERT(st == DOAPP_OK);
  have_u8_grid = false;
}

void PlanningState::updateOccupancyGrid(const doapp_host::OccupancyGridMsg &msg) {
  synchronize();
  doapp_status st =
      doapp_grid_from_msg(ctx, msg.size, msg.granularity, msg.grid.data(), msg.grid.size());
  DOAPP_ASSERT(st == DOAPP_OK);
  have_u8_grid = true;
}

void PlanningState::updateDepthFrames(const uint16_t *depth, unsigned int num_cams,
                                      unsigned int width, unsigned int height,
                                      const double *intrinsics, const float *extrinsics,
                                      double crop_size, bool clear_first) {
  synchronize();
  if (!have_u8_grid) {
    doapp_status st = doapp_grid_alloc(ctx, DOAPP_GRID_U8, DOAPP_FRAME_CDG, (uint32_t)dims.length,
                                       (uint32_t)dims.width, (uint32_t)dims.height,
                                       dims.resolution);
    DOAPP_ASSERT(st == DOAPP_OK);
    have_u8_grid = true;
  }
  doapp_status st = doapp_grid_from_depth(ctx, depth, num_cams, width, height, intrinsics,
                                          extrinsics, crop_size, 1, clear_first ? 1 : 0, nullptr);
  DOAPP_ASSERT(st == DOAPP_OK);
}

void PlanningState::buildDistanceField() {
  synchronize();
  DOAPP_ASSERT(have_u8_grid);
  doapp_status st = doapp_grid_distance_field(ctx);
  DOAPP_ASSERT(st == DOAPP_OK);
  have_u8_grid = false; // the grid now holds float distances
}

void PlanningState::useUrdfKinematics(bool on) {
  synchronize();
  doapp_fk_model model;
  doapp_fk_model_rexarm(&model);
  doapp_status st = doapp_set_fk_model(ctx, on ? &model : nullptr);
  DOAPP_ASSERT(st == DOAPP_OK);
}

void PlanningState::updatePointCloud(const doapp_host::PointCloudXYZ &cloud, bool clear_first) {
  synchronize();
  if (!have_u8_grid) {
    doapp_status st = doapp_grid_alloc(ctx, DOAPP_GRID_U8, DOAPP_FRAME_CDG, (uint32_t)dims.length,
                                       (uint32_t)dims.width, (uint32_t)dims.height,
                                       dims.resolution);
    DOAPP_ASSERT(st == DOAPP_OK);
    have_u8_grid = true;
  }
  doapp_status st =
      doapp_grid_scatter(ctx, cloud.xyz.data(), cloud.xyz.size() / 3, 1, clear_first ? 1 : 0);
  DOAPP_ASSERT(st == DOAPP_OK);
}
