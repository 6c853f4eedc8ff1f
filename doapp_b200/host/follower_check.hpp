// follower_check.hpp — the consumer side of the planner's output, replayed without servos or
// ROS: what dynamixel_driver/src/trajectory_follower.cpp:131-215 does with a published
// JointTrajectory.  The follower keeps two indices into the points,
// prev_waypt_ / next_waypt_ (:139-144), and on every 100 Hz tick
//   * interpolates the reported position between the two points by time (:164-179),
//   * commands the servos with points[prev_waypt_].positions / .velocities, the first
//     joints_.num_joints (6) entries of each (:183-199),
//   * advances both indices once `now` has passed the segment's end (:206-211).
// As shipped, the segment end is read from points[prev_waypt_] too (:159-160), so the segment
// has zero length: the reported position jumps to the next point (progress = min(1, x / 0))
// and the indices advance as soon as `now` passes the PREVIOUS point's time, one segment
// early; `intended` reads the end from points[next_waypt_].  The replay does both, plus the
// checks a drop-in planner has to satisfy for this consumer: every point carries at least the
// servo count of entries, time_from_start does not decrease, and the velocity a point
// announces is the finite difference towards the next point (planning_state.cu:186,196-198
// write look-forward velocities), i.e. positions, velocities and times are consistent.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <vector>

#include "ros_mirror.hpp"

namespace doapp_host {

struct FollowerReplay {
  bool well_formed = false;      // >= num_joints entries per point, times non-decreasing
  std::size_t ticks = 0;         // ticks until the last point became the commanded one
  double duration = 0.0;         // seconds of those ticks
  double max_command_step = 0.0; // largest |commanded(t+1) - commanded(t)| over joints, radians
  double max_interp_gap = 0.0;   // largest |interpolated report - straight line| (sanity: 0)
  double max_lead = -1e300;      // max over ticks of (stamp of the commanded point - now), seconds
  double max_velocity_mismatch = 0.0; // max |(p[i+1] - p[i]) / (t[i+1] - t[i]) - v[i]|, planner joints
  std::vector<std::vector<double>> commanded; // one row per tick, num_joints entries
};

// tick_s: the follower's loop period (100 Hz: trajectory_follower_node.cpp).  planner_joints:
// how many leading entries the planner optimises (the rest are padding held constant).
inline FollowerReplay replay_follower(const JointTrajectory &msg, std::size_t num_joints,
                                      std::size_t planner_joints, double tick_s, bool intended,
                                      std::size_t max_ticks = 100000) {
  FollowerReplay r;
  const auto &pts = msg.points;
  if (pts.empty()) return r;
  r.well_formed = true;
  for (std::size_t i = 0; i < pts.size(); ++i) {
    if (pts[i].positions.size() < num_joints || pts[i].velocities.size() < num_joints)
      r.well_formed = false; // :176-178,194-199 would read past the end
    if (i && pts[i].time_from_start < pts[i - 1].time_from_start) r.well_formed = false;
  }
  if (!r.well_formed) return r;
  // consistency of the message itself (planning_state.cu:179-209 writes finite differences)
  for (std::size_t i = 0; i + 2 < pts.size(); ++i) { // the goal point is appended, not planned
    const double dt = pts[i + 1].time_from_start - pts[i].time_from_start;
    if (!(dt > 0.0)) continue; // points 0 and 1 share stamp 0 (planning_state.cu:190,200)
    for (std::size_t j = 0; j < planner_joints; ++j) {
      const double fd = (pts[i + 1].positions[j] - pts[i].positions[j]) / dt;
      r.max_velocity_mismatch = std::max(r.max_velocity_mismatch, std::fabs(fd - pts[i].velocities[j]));
    }
  }
  std::size_t prev = 0, next = pts.size() > 1 ? 1 : 0; // :139-144
  double now = 0.0;                                     // header.stamp is the origin
  std::vector<double> last;
  while (r.ticks < max_ticks) {
    const double prev_time = pts[prev].time_from_start;                       // :157-158
    const double next_time = pts[intended ? next : prev].time_from_start;     // :159-160
    const double segment = next_time - prev_time;
    std::vector<double> report(num_joints);
    if (prev == next) {
      for (std::size_t j = 0; j < num_joints; ++j) report[j] = pts.back().positions[j]; // :165
    } else {
      double progress = segment > 0.0 ? std::min(1.0, (now - prev_time) / segment) : 1.0; // :169-170
      progress = std::max(0.0, progress);
      for (std::size_t j = 0; j < num_joints; ++j) { // :176-179
        report[j] = (1.0 - progress) * pts[prev].positions[j] + progress * pts[next].positions[j];
        const double lo = std::min(pts[prev].positions[j], pts[next].positions[j]);
        const double hi = std::max(pts[prev].positions[j], pts[next].positions[j]);
        r.max_interp_gap = std::max(r.max_interp_gap, std::max(lo - report[j], report[j] - hi));
      }
    }
    std::vector<double> cmd(pts[prev].positions.begin(), pts[prev].positions.begin() + num_joints); // :187-188
    if (!last.empty())
      for (std::size_t j = 0; j < num_joints; ++j)
        r.max_command_step = std::max(r.max_command_step, std::fabs(cmd[j] - last[j]));
    last = cmd;
    r.commanded.push_back(std::move(cmd));
    r.max_lead = std::max(r.max_lead, prev_time - now);
    r.ticks += 1;
    if (prev + 1 == pts.size()) break; // the goal point is being commanded
    if (now > next_time) { // :206-211
      prev = next;
      if (next + 1 < pts.size()) next++;
    }
    now += tick_s;
  }
  r.duration = now;
  return r;
}

} // namespace doapp_host
