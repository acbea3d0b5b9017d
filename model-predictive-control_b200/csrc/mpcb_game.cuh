// mpcb_game.cuh — the game-theoretic lane-change decision (reference game_theory.py:115-244) for a batch of scenarios.
//
// What it replaces: `Car.get_total_payoff(cars, target_lane)` evaluated for target lanes 1 and 2, the decision of the
// reference's sweep (game_theory.py:377-379: change when the lane-2 payoff is larger) and `get_bezier_control_points`
// towards the car in front (game_theory.py:92-113), whose control points become the MPC's Bezier road (config 4).
// One lane per scenario: the work is a few dozen branches and two transcendentals per pair of cars, so the kernel is
// bound by its loads — ego (3 doubles) and M car records (3 doubles each) read once, 15 doubles + 1 int written.
// Records are (x, v, lane); lane <= 0 marks an empty slot of a padded list.
#pragma once
#include "mpcb_device.cuh"

namespace mpcb {

struct GamePar { double L, W, l, th, tlc, td, ti, tau, a_max, h, Lf; };  // game_theory.py:23-56

struct GCar { double x, v; int lane; };

// game_theory.py:115-153, q1 = 0.65, q2 = 0.35
MPCB_DEV double g_safety_distance(const GamePar& c, const GCar& s, const GCar& o, int target, double sin_th) {
  const double q1 = 0.65, q2 = 0.35;
  const double dv = s.v - o.v;
  const double react = q2 * (dv * c.tau + c.ti / 2 + dv * dv / (2 * c.a_max));
  if (s.lane == o.lane) {
    if (s.x > o.x) return fabs(s.x - o.x);
    if (target == s.lane) return q1 * s.v + c.td + react + c.l;
    if (s.v > o.v) return s.v - o.v * c.tlc / 2 + c.L + c.W / 2 * sin_th;
    return q1 * s.v * c.td + c.l;
  }
  if (s.x < o.x) {
    if (s.v > o.v) return s.v - o.v * c.tlc / 2 + c.L - c.W / 2 * sin_th + q1 * s.v * c.td + react;
    return q1 * s.v * c.td + c.l;
  }
  if (s.v < o.v) {
    const double du = o.v - s.v;
    return du * 3 / 4 * c.tlc + c.L + q1 * o.v * c.td + q2 * (du * c.tau + c.ti / 2 + du * du / (2 * c.a_max));
  }
  return q1 * o.v * c.td + c.l;
}

// Everything a car's view of the scene needs: the list, one index left out, and an optional extra car.
struct GScene {
  const double* cars;  // [M, 3]
  int M, skip;
  bool has_extra;
  GCar extra;
  MPCB_DEV int count() const { return M + (has_extra ? 1 : 0); }
  MPCB_DEV bool get(int k, GCar& o) const {
    if (k == skip) return false;
    if (k < M) { o.x = cars[3 * k]; o.v = cars[3 * k + 1]; o.lane = (int)cars[3 * k + 2]; }
    else o = extra;
    return o.lane > 0;
  }
};

// game_theory.py:155-176
MPCB_DEV double g_safety_payoff(const GamePar& c, const GCar& s, const GScene& sc, int target, double sin_th) {
  double payoff = 1.0, temp = 1.0;
  for (int k = 0; k < sc.count(); ++k) {
    GCar o;
    if (!sc.get(k, o)) continue;
    if (s.lane != o.lane && s.lane == target) continue;
    const double Sk = g_safety_distance(c, s, o, target, sin_th), Dk = fabs(s.x - o.x);
    if (Dk >= fabs(Sk)) temp = 1.0;
    if (Dk <= c.l) temp = -1.0;
    if (c.l < Dk && Dk < fabs(Sk)) temp = log(Dk / Sk + 1.0) / log(2.0);
    if (temp < payoff) payoff = temp;
  }
  return payoff;
}

// game_theory.py:58-75 + 178-189
MPCB_DEV bool g_front(const GCar& s, const GScene& sc, int lane, GCar& f) {
  bool any = false;
  for (int k = 0; k < sc.count(); ++k) {
    GCar o;
    if (!sc.get(k, o) || o.lane != lane || !(o.x > s.x)) continue;
    if (!any || f.x > o.x) { f = o; any = true; }
  }
  return any;
}
MPCB_DEV double g_velocity_payoff(const GCar& s, const GScene& sc, int target) {
  GCar f;
  if (!g_front(s, sc, target, f)) return 1.0;
  if (f.v == 0.0) return -1.0;
  if (f.v >= 2.0 * s.v) return 1.0;
  return (f.v - s.v) / s.v;
}

// game_theory.py:205-244, a = 0.6, b = 0.4
MPCB_DEV double g_total_payoff(const GamePar& c, const GCar& ego, const double* cars, int M, int target, double sin_th) {
  const double a = 0.6, b = 0.4;
  GScene all{cars, M, -1, false, GCar{0.0, 0.0, 0}};
  const double total = a * g_safety_payoff(c, ego, all, target, sin_th) + b * g_velocity_payoff(ego, all, target);
  int behind = -1;  // the nearest car behind ego in lane 2 (game_theory.py:77-90)
  GCar cb{0.0, 0.0, 0};
  for (int k = 0; k < M; ++k) {
    GCar o;
    if (!all.get(k, o) || o.lane != 2 || !(o.x < ego.x)) continue;
    if (behind < 0 || cb.x < o.x) { behind = k; cb = o; }
  }
  double total_behind = 0.0;
  if (behind >= 0) {  // its own payoff for keeping lane 2, with a copy of ego in lane 2 when that is ego's target
    GScene rest{cars, M, behind, target == 2, GCar{ego.x, ego.v, 2}};
    total_behind = a * g_safety_payoff(c, cb, rest, 2, sin_th) + b * g_velocity_payoff(cb, rest, 2);
  }
  return total + total_behind;
}

__global__ void k_game_payoff(int64_t B, int M, GamePar c, const double* __restrict__ ego, const double* __restrict__ cars,
                              double* __restrict__ payoff, int32_t* __restrict__ decision, double* __restrict__ ctrl,
                              double* __restrict__ tca) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const GCar e{ego[3 * b], ego[3 * b + 1], (int)ego[3 * b + 2]};
  const double* cs = cars + b * M * 3;
  const double sin_th = sin(c.th);
  const double p1 = g_total_payoff(c, e, cs, M, 1, sin_th), p2 = g_total_payoff(c, e, cs, M, 2, sin_th);
  payoff[2 * b] = p1;
  payoff[2 * b + 1] = p2;
  if (decision) decision[b] = p2 > p1 ? 1 : 0;  // game_theory.py:377-379
  if (ctrl) {  // game_theory.py:92-113 towards the car in front in ego's own lane, i = 5
    GScene all{cs, M, -1, false, GCar{0.0, 0.0, 0}};
    GCar f;
    double* P = ctrl + 12 * b;
    if (g_front(e, all, e.lane, f) && e.v != f.v) {
      const double i = 5.0;
      const double Di = (c.Lf + c.l) * cos(atan2(c.W, 2.0 * c.Lf) - c.th);
      const double D1 = f.x - e.x;
      const double tc1 = D1 / (e.v - f.v);
      const double Px2 = e.v * tc1 - Di, Px5 = 2.0 * Px2;
      P[0] = 0.0; P[1] = (Px2 - 0.0) / i; P[2] = Px2; P[3] = Px2; P[4] = Px5 - (Px5 - Px2) / i; P[5] = Px5;
      P[6] = 0.0; P[7] = 0.0; P[8] = 0.0; P[9] = c.h; P[10] = c.h; P[11] = c.h;
      if (tca) tca[b] = Px2 / (e.v - f.v);
    } else {
      const double nn = nan("");
      for (int q = 0; q < 12; ++q) P[q] = nn;
      if (tca) tca[b] = nn;
    }
  }
}

}  // namespace mpcb
