// motion_planner_main.cpp — the planner node (executable `pertubator_main`, node
// "motion_planner") of perturbator/src/main.cu:14-45 on top of the re-hosted PlanningState.
//
// With -DDOAPP_WITH_ROS this is the ROS node: same node name, same topics (sub joint_state and
// goal_state, queue 1; pub /joint_trajectory, queue 2), same grid dimensions and problem shape
// (main.cu:18-21), same 1 Hz plan / sleep / publish cycle (main.cu:29-44).
// Without ROS (this image has none) the same cycle is driven from a scripted sequence of
// joint_state / goal_state messages (the goals of task/src/main.cpp:33-42) through the mirror
// types of ros_mirror.hpp, so the loop is compiled and exercised here as well.
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <memory>

#include "planning_state.hpp"

namespace {
constexpr unsigned int kNumJoints = 4; // doapp::num_joints, perturbator/include/common.cuh:10
}

#ifdef DOAPP_WITH_ROS
#include "trajectory_msgs/JointTrajectory.h"

int main(int argc, char **argv) {
  ros::init(argc, argv, "motion_planner");
  ros::NodeHandle node_handle;
  const doapp::cpu_distance_grid::Dimensions grid_dims = {100, 100, 100, 0.01f}; // main.cu:18
  const unsigned int k = 1000, n = 50, d = kNumJoints, m = 1024 / n / d;         // main.cu:20-21
  std::cout << "Running with k = " << k << ", n = " << n << ", m = " << m << std::endl;
  PlanningState planning_state(k, m, n, d, grid_dims);
  ros::Subscriber state_subscriber =
      node_handle.subscribe("joint_state", 1, &PlanningState::updateArmState, &planning_state);
  ros::Subscriber goal_subscriber =
      node_handle.subscribe("goal_state", 1, &PlanningState::updateGoalState, &planning_state);
  ros::Rate loop_rate(1);
  ros::Publisher trajectory_publisher =
      node_handle.advertise<trajectory_msgs::JointTrajectory>("/joint_trajectory", 2);
  while (ros::ok()) {
    ros::spinOnce();
    if (planning_state.have_goal()) planning_state.plan(); // asynchronous (planning_state.cu:79-84)
    loop_rate.sleep();                                     // the optimiser's time budget
    if (planning_state.have_goal()) planning_state.publish_trajectory(trajectory_publisher);
  }
  return 0;
}
#else
int main(int argc, char **argv) {
  const unsigned int cycles = argc > 1 ? (unsigned int)std::atoi(argv[1]) : 3;
  const doapp::cpu_distance_grid::Dimensions grid_dims = {100, 100, 100, 0.01f}; // main.cu:18
  const unsigned int k = 1000, n = 50, d = kNumJoints, m = 1024 / n / d;         // main.cu:20-21
  std::cout << "Running with k = " << k << ", n = " << n << ", m = " << m << std::endl;
  PlanningState planning_state(k, m, n, d, grid_dims);
  planning_state.setIterations(argc > 2 ? (unsigned int)std::atoi(argv[2]) : 20);
  // goals of task/src/main.cpp:33-42 (4 joints)
  const double goals[3][4] = {{0.0, 1.0, 0.0, 0.5}, {0.0, -1.0, 1.5, -1.5}, {1.0, 0.5, -0.5, 0.0}};
  doapp_host::Publisher trajectory_publisher;
  auto state = std::make_shared<doapp_host::JointState>();
  state->position = {0.0, 0.0, 0.0, 0.0};
  for (unsigned int cycle = 0; cycle < cycles; ++cycle) {
    // "spinOnce": the callbacks the subscriptions would deliver
    planning_state.updateArmState(state);
    auto goal = std::make_shared<doapp_host::JointState>();
    goal->position.assign(goals[cycle % 3], goals[cycle % 3] + 4);
    planning_state.updateGoalState(goal);
    if (planning_state.have_goal()) planning_state.plan();
    // (the node sleeps one loop period here; the bounded optimiser needs no wall-clock budget)
    if (planning_state.have_goal()) planning_state.publish_trajectory(trajectory_publisher);
    const doapp_host::JointTrajectory &msg = trajectory_publisher.published.back();
    std::printf("cycle %u: %zu points, cost %.6g, ends at", cycle, msg.points.size(),
                planning_state.last_cost());
    for (double q : msg.points.back().positions) std::printf(" %.3f", q);
    std::printf("\n");
    state->position = msg.points.back().positions; // the arm follows the trajectory
  }
  return trajectory_publisher.published.size() == cycles ? 0 : 1;
}
#endif
