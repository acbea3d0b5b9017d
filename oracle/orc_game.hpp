// orc_game.hpp — CPU ORACLE (test infrastructure, NOT product code): the game-theoretic lane-change decision of
// reference game_theory.py:115-244, restated for records (x, v, lane) instead of Car objects.  Pinned to golden vectors
// produced by the reference's own code (tests/golden/gen_golden_game.py -> reference_game_theory.json).
#pragma once
#include <cmath>

namespace orc {

// Car constants in game_theory.py:23-56 order of use: L, W, l, theta_max, tlc, td, ti, tau, a_max, h, Lf.
struct CarPar { double L, W, l, th, tlc, td, ti, tau, a_max, h, Lf; };
inline CarPar car_defaults() { return {4.2, 1.8, 3.0, 3.2 / 180 * M_PI, 5.17, 1.2, 0.15, 0.9, 7.0, 3.75, 1.0}; }

struct CarRec { double x, v; int lane; };  // lane <= 0: empty slot of a padded list (skipped everywhere)

// game_theory.py:115-153 (q1 = 0.65, q2 = 0.35); every car carries the same constants, so car.a_max == self.a_max.
inline double safety_distance(const CarPar& c, const CarRec& s, const CarRec& o, int target_lane) {
  const double q1 = 0.65, q2 = 0.35;
  if (s.lane == o.lane) {
    if (s.x > o.x) return std::fabs(s.x - o.x);
    if (target_lane == s.lane)
      return q1 * s.v + c.td + q2 * ((s.v - o.v) * c.tau + c.ti / 2 + (s.v - o.v) * (s.v - o.v) / (2 * c.a_max)) + c.l;
    if (s.v > o.v) return s.v - o.v * c.tlc / 2 + c.L + c.W / 2 * std::sin(c.th);  // S01
    return q1 * s.v * c.td + c.l;
  }
  if (s.x < o.x) {  // S02
    if (s.v > o.v)
      return s.v - o.v * c.tlc / 2 + c.L - c.W / 2 * std::sin(c.th) + q1 * s.v * c.td +
             q2 * ((s.v - o.v) * c.tau + c.ti / 2 + (s.v - o.v) * (s.v - o.v) / (2 * c.a_max));
    return q1 * s.v * c.td + c.l;
  }
  if (s.v < o.v)  // S03
    return (o.v - s.v) * 3 / 4 * c.tlc + c.L + q1 * o.v * c.td +
           q2 * ((o.v - s.v) * c.tau + c.ti / 2 + (o.v - s.v) * (o.v - s.v) / (2 * c.a_max));
  return q1 * o.v * c.td + c.l;
}

// game_theory.py:155-176.  `skip` = index of a car left out (the car behind never sees itself); `extra` (nullable) is
// appended to the list (the phantom copy of ego in lane 2, game_theory.py:225-229).
inline double safety_payoff(const CarPar& c, const CarRec& s, const CarRec* cars, int M, int skip, const CarRec* extra,
                            int target_lane) {
  double payoff = 1, temp = 1;
  for (int k = 0; k < M + (extra ? 1 : 0); ++k) {
    if (k == skip) continue;
    const CarRec& o = k < M ? cars[k] : *extra;
    if (o.lane <= 0) continue;
    if (s.lane != o.lane && s.lane == target_lane) continue;
    const double Sk = safety_distance(c, s, o, target_lane), Dk = std::fabs(s.x - o.x);
    if (Dk >= std::fabs(Sk)) temp = 1;
    if (Dk <= c.l) temp = -1;
    if (c.l < Dk && Dk < std::fabs(Sk)) temp = std::log(Dk / Sk + 1) / std::log(2.0);
    if (temp < payoff) payoff = temp;
  }
  return payoff;
}

// game_theory.py:58-75: the nearest car ahead in `lane` (first one wins ties); -1 if none, M = the extra car.
inline int car_in_front(const CarRec& s, const CarRec* cars, int M, int skip, const CarRec* extra, int lane) {
  int best = -1;
  double bx = 0;
  for (int k = 0; k < M + (extra ? 1 : 0); ++k) {
    if (k == skip) continue;
    const CarRec& o = k < M ? cars[k] : *extra;
    if (o.lane <= 0 || o.lane != lane || !(o.x > s.x)) continue;
    if (best < 0 || bx > o.x) { best = k; bx = o.x; }
  }
  return best;
}

// game_theory.py:178-189
inline double velocity_payoff(const CarRec& s, const CarRec* cars, int M, int skip, const CarRec* extra, int target_lane) {
  const int f = car_in_front(s, cars, M, skip, extra, target_lane);
  if (f < 0) return 1;
  const CarRec& o = f < M ? cars[f] : *extra;
  if (o.v == 0) return -1;
  if (o.v >= 2 * s.v) return 1;
  return (o.v - s.v) / s.v;
}

// game_theory.py:205-244 (a = 0.6, b = 0.4); the car behind is always looked for in lane 2 (:77-90).
inline double total_payoff(const CarPar& c, const CarRec& ego, const CarRec* cars, int M, int target_lane) {
  const double a = 0.6, b = 0.4;
  double total = a * safety_payoff(c, ego, cars, M, -1, nullptr, target_lane) +
                 b * velocity_payoff(ego, cars, M, -1, nullptr, target_lane);
  int behind = -1;
  double bx = 0;
  for (int k = 0; k < M; ++k) {
    if (cars[k].lane != 2 || !(cars[k].x < ego.x)) continue;
    if (behind < 0 || bx < cars[k].x) { behind = k; bx = cars[k].x; }
  }
  double total_behind = 0;
  if (behind >= 0) {
    const CarRec phantom{ego.x, ego.v, 2};
    const CarRec* extra = target_lane == 2 ? &phantom : nullptr;
    total_behind = a * safety_payoff(c, cars[behind], cars, M, behind, extra, 2) +
                   b * velocity_payoff(cars[behind], cars, M, behind, extra, 2);
  }
  return total + total_behind;
}

// game_theory.py:92-113 towards `front` (i = 5): P[0..5] = Px, P[6..11] = Py (relative to ego, lane centre 0 -> h).
inline void lane_change_points(const CarPar& c, const CarRec& s, const CarRec& front, double i, double* P, double* tca) {
  const double Li = c.Lf + c.l;
  const double Di = Li * std::cos(std::atan2(c.W, 2 * c.Lf) - c.th);
  const double D1 = front.x - s.x;
  const double tc1 = D1 / (s.v - front.v);
  const double Px2 = s.v * tc1 - Di, Px5 = 2 * Px2;
  P[0] = 0; P[1] = (Px2 - 0) / i; P[2] = Px2; P[3] = Px2; P[4] = Px5 - (Px5 - Px2) / i; P[5] = Px5;
  P[6] = P[7] = P[8] = 0; P[9] = P[10] = P[11] = c.h;
  *tca = Px2 / (s.v - front.v);
}

}  // namespace orc
