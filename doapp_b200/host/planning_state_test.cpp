// planning_state_test.cpp — drives PlanningState the way perturbator/src/main.cu:24-43 does
// (joint_state + goal_state callbacks, plan(), publish_trajectory()) without ROS, and prints the
// published JointTrajectory as JSON for tests/test_host_gpu.py to check against the oracle.
//   planning_state_test k m n d grid_n resolution iterations cloud.bin|- [follower_joints [edt|kdtree|- [urdf]]]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#include "follower_check.hpp"
#include "planning_state.hpp"

int main(int argc, char **argv) {
  if (argc < 9) {
    std::fprintf(stderr, "usage: %s k m n d grid_n resolution iterations cloud.bin|-\n", argv[0]);
    return 2;
  }
  const unsigned k = std::atoi(argv[1]), m = std::atoi(argv[2]), n = std::atoi(argv[3]),
                 d = std::atoi(argv[4]), gn = std::atoi(argv[5]), iters = std::atoi(argv[7]);
  const double res = std::atof(argv[6]);
  const doapp::cpu_distance_grid::Dimensions dims = {gn, gn, gn, res};
  PlanningState planning_state(k, m, n, d, dims);
  planning_state.setIterations(iters);
  if (argc > 9) planning_state.setFollowerJoints(std::atoi(argv[9]));

  if (argv[8][0] != '-') { // the grid feed: an xyz cloud scattered into the grid
    FILE *f = std::fopen(argv[8], "rb");
    if (!f) return 3;
    std::fseek(f, 0, SEEK_END);
    long bytes = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    doapp_host::PointCloudXYZ cloud;
    cloud.xyz.resize(bytes / sizeof(float));
    if (std::fread(cloud.xyz.data(), 1, bytes, f) != (size_t)bytes) return 3;
    std::fclose(f);
    const bool want_tree = argc > 10 && std::strcmp(argv[10], "kdtree") == 0;
    if (want_tree) {
#ifdef DOAPP_WITH_NANOFLANN
      // the reference's own entry point: a nanoflann KD-tree over the obstacle points
      // (planning_state.cuh:25-28)
      const doapp::cpu_distance_grid::MatrixPointcloudAdaptor adaptor(cloud.xyz);
      doapp::cpu_distance_grid::KDTree tree(3, adaptor, nanoflann::KDTreeSingleIndexAdaptorParams(10));
      tree.buildIndex();
      planning_state.updateDistanceGrid(tree);
#else
      return 7; // built without nanoflann
#endif
    } else {
      planning_state.updatePointCloud(cloud);
      if (argc > 10 && argv[10][0] != '-')
        planning_state.buildDistanceField(); // sphere test on the derived distance grid
    }
  }
  if (argc > 11) planning_state.useUrdfKinematics(true);

  auto state = std::make_shared<doapp_host::JointState>();
  state->position = {0.0, 0.0, 0.0, 0.0, 0.0, 0.7}; // 6 servos reported, truncated to d
  auto goal = std::make_shared<doapp_host::JointState>();
  goal->position = {0.0, -1.0, 1.5, -1.5, 0.0}; // task/src/main.cpp:39 + wrist_rotate
  planning_state.updateArmState(state);
  planning_state.updateGoalState(goal);
  doapp_host::Publisher pub;
  if (!planning_state.have_goal()) return 4;
  planning_state.plan();
  planning_state.publish_trajectory(pub); // main.cu:40-43
  if (planning_state.have_goal() || pub.published.size() != 1) return 5;
  planning_state.updateGoalState(goal); // unchanged goal: no replanning (planning_state.cuh:39)
  if (planning_state.have_goal()) return 6;

  const doapp_host::JointTrajectory &msg = pub.published[0];
  std::printf("{\"cost\": %.9g, \"points\": [", planning_state.last_cost());
  for (size_t i = 0; i < msg.points.size(); ++i) {
    const auto &p = msg.points[i];
    std::printf("%s{\"t\": %.9g, \"positions\": [", i ? ", " : "", p.time_from_start);
    for (size_t j = 0; j < p.positions.size(); ++j)
      std::printf("%s%.9g", j ? ", " : "", p.positions[j]);
    std::printf("], \"velocities\": [");
    for (size_t j = 0; j < p.velocities.size(); ++j)
      std::printf("%s%.9g", j ? ", " : "", p.velocities[j]);
    std::printf("]}");
  }
  std::printf("]");
  // the consumer's side: replay trajectory_follower.cpp:148-215 at its 100 Hz tick
  const size_t servos = argc > 9 ? (size_t)std::atoi(argv[9]) : d;
  for (int intended = 0; intended < 2; ++intended) {
    const doapp_host::FollowerReplay r =
        doapp_host::replay_follower(msg, servos > d ? servos : d, d, 0.01, intended != 0);
    std::printf(", \"follower_%s\": {\"well_formed\": %d, \"ticks\": %zu, \"duration\": %.9g, "
                "\"max_command_step\": %.9g, \"max_interp_gap\": %.9g, "
                "\"max_velocity_mismatch\": %.9g, \"max_lead\": %.9g, \"last_command\": [",
                intended ? "intended" : "as_shipped", r.well_formed ? 1 : 0, r.ticks, r.duration,
                r.max_command_step, r.max_interp_gap, r.max_velocity_mismatch, r.max_lead);
    if (!r.commanded.empty())
      for (size_t j = 0; j < r.commanded.back().size(); ++j)
        std::printf("%s%.9g", j ? ", " : "", r.commanded.back()[j]);
    std::printf("]}");
  }
  std::printf("}\n");
  return 0;
}
