// Host optimizer: turns the reduced [error, gradient] of an epoch into the next parameter vector.
//
// Behavioural contract: the numbers Optimizer::update_network of the reference produces, call for call
// (Optimizer.cpp:884-1064 with Rprop :521-593, steepest descent :400-449 / :457-491, L-BFGS :607-696, the Brent-style
// line minimisation :709-869 and the perturbation scheme :915-938, 1008-1043) -- tests/golden/opt_*.npz hold recorded
// call sequences of the reference and tests/test_host_layer.py replays them bit for bit.  The organisation is this
// repository's own: the trajectory lives in flat fixed-capacity rings (no deque of vectors), every algorithm is a step
// function over an explicit state record, and the line search is a small state machine (Fresh / Expanding / Bracketed)
// instead of deques whose length encodes the state.  It runs on the host on the already reduced gradient (P <= ~50k
// doubles): every rank holds the identical vector, takes the identical step, and the reference's broadcast of the new
// parameters (Optimizer.cpp:957) has nothing left to do.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <limits>

#include "pph.h"

namespace pph {

namespace {

inline int sgn(REAL v) { return (REAL(0) < v) - (v < REAL(0)); }

// plain left-to-right sum of products (the translation unit is built without FMA contraction)
REAL inner(const REAL* a, const REAL* b, int n)
{
  REAL r = 0.0;
  for (int i = 0; i < n; i++) r += a[i] * b[i];
  return r;
}
inline REAL inner(const std::vector<REAL>& a, const std::vector<REAL>& b) { return inner(a.data(), b.data(), (int)a.size()); }
inline REAL length(const REAL* a, int n) { return std::sqrt(inner(a, a, n)); }

// v / |v|, element by element
void unit_vector(const REAL* v, int n, std::vector<REAL>& out)
{
  const REAL len = length(v, n);
  out.resize(n);
  for (int i = 0; i < n; i++) out[i] = v[i] / len;
}

// Gaussian deviate from two libc rand() draws, first under the logarithm, second in the cosine (RAND::Normal, Common.h:72-75)
REAL gaussian_from_rand(REAL mean, REAL sd)
{
  const REAL u_log = (REAL)rand() / RAND_MAX;
  const REAL u_cos = (REAL)rand() / RAND_MAX;
  return mean + sd * std::sqrt(-2.0 * std::log(u_log)) * std::cos(2.0 * 3.141592653589793 * u_cos);
}

const int kKeep = 20;      // points kept of the trajectory (Optimizer.cpp:962-966)

}  // namespace

// ---- Ring: fixed-capacity FIFO of equal-length rows, one flat allocation -------------------------------------------------
void Optimizer::Ring::reset(int width, int capacity)
{
  w = width; cap = capacity; first = 0; n = 0;
  buf.assign((size_t)std::max(w, 1) * cap, 0.0);
}
REAL* Optimizer::Ring::append()
{
  if (n == cap) drop_oldest();
  REAL* p = &buf[(size_t)((first + n) % cap) * w];
  n++;
  return p;
}
void Optimizer::Ring::drop_oldest() { first = (first + 1) % cap; n--; }
const REAL* Optimizer::Ring::oldest(int i) const { return &buf[(size_t)((first + i) % cap) * w]; }
const REAL* Optimizer::Ring::newest(int age) const { return oldest(n - 1 - age); }
REAL* Optimizer::Ring::newest_mut() { return &buf[(size_t)((first + n - 1) % cap) * w]; }

// ---- construction -------------------------------------------------------------------------------------------------------
Optimizer::Optimizer(Parallel* mpi_in, const FunctionalParams& params, int Nsystems) : do_print(true), mpi(mpi_in), F(params)
{
  n_structures = Nsystems;
  n_par = 0;
  epoch = 1;
  done = want_checkpoint = want_backup = false;
  line_done = true;
  report_validation = false;
  val_sse = 0;
  val_count = 0;
  // Optimizer::set_training_algorithm (Optimizer.cpp:139-166)
  const std::string& name = params.training_algorithm;
  may_switch = false;
  if (name == "default") { method = kRprop; may_switch = true; }
  else if (name == "rprop") method = kRprop;
  else if (name == "steepest_descent") method = kDescent;
  else if (name == "steepest_descent_with_line_opt") method = kDescentLine;
  else if (name == "bfgs") method = kLbfgs;
  else PPH_ERROR("Training algorithm \"" + name + "\" is not implemented");
  first_size = 1e-4;
  size_ceiling = 0.5;
  size_floor = 1000.0 * std::numeric_limits<REAL>::epsilon();
  restart_size = first_size;
  step_len = first_size;
  rp_started = false;
  dir_set = false;
  lbfgs_memory = 77;
  ls.phase = LineSearch::kFresh;
  ls.npts = 0;
  ls.ok_E = ls.ok_a = ls.ok_s = ls.c_a = ls.c_s = ls.d_a = 0.0;
  ls.bisected = false;
  ls.resolution = 100 * std::numeric_limits<REAL>::epsilon();
  // perturbation scheme: the accessor of the reference returns int (Functional_params.h:198-201), so amplitudes below 1 switch it off
  pt.ceiling = (int)F.lambda_max;
  pt.amplitude_now = pt.ceiling;
  pt.enabled = true;
  pt.alpha = pt.dphi = pt.phase = pt.rate = 0.0;
  pt.cycles = 0;
  pt.t_on = 0;
  pt.best_E = 9e9;
  pt.best_norm = 0.0;
  pt.floor_E = 0.0;
}

void Optimizer::init(const std::vector<REAL>& x0)
{
  // table header (Optimizer::internal_init, Optimizer.cpp:242-289)
  if (mpi->io_node() && do_print) {
    if (report_validation) {
      std::cout << "Iteration    Ɛ_train = Σ(E²)    Ɛ_val = Σ(E²)       |∇Ɛ|" << std::endl;
      std::cout << "-------------------------------------------------------------" << std::endl;
    } else {
      std::cout << "Iteration    Ɛ = Σ(E²)       |∇Ɛ|" << std::endl;
      std::cout << "-------------------------------------" << std::endl;
    }
  }
  n_par = (int)x0.size();
  points.reset(n_par, kKeep + 2);
  slopes.reset(n_par, kKeep + 2);
  errors.reset(1, kKeep + 2);
  std::copy(x0.begin(), x0.end(), points.append());
  current = x0;
  rp_size.assign(n_par, first_size);
  rp_move.assign(n_par, 0.0);
  rp_live.assign(n_par, 0);
  rp_started = false;
  dir_set = false;
  dir.clear();
}

// ---- Rprop with secant steps (Optimizer.cpp:521-593) -----------------------------------------------------------------------
// Per parameter: a live component whose gradient kept its sign grows its step (x 1.2, capped except for the very last
// parameter); a sign change parks the component for one call; a parked component (or a vanishing product) takes a secant
// step from the gradients two calls apart, clamped, and goes live again.  The very first call moves every parameter by
// -first_size whatever the gradient says.
std::vector<REAL> Optimizer::step_rprop()
{
  line_done = true;
  const int have = slopes.size();
  std::vector<REAL> next(points.newest(0), points.newest(0) + n_par);
  const REAL* g = slopes.newest(0);
  if (may_switch && epoch % 67 == 0 && have > 1 && inner(g, g, n_par) < 1e11 && rp_started) {
    // default schedule: once every component's secant slope is non-negative the surface looks convex -> steepest descent
    const REAL* g_prev = slopes.newest(1);
    bool convex = n_par > 0;
    for (int i = 0; i < n_par && convex; i++)
      if ((g[i] - g_prev[i]) / rp_move[i] < 0) convex = false;
    if (convex) {
      for (int k = 0; k < n_par; k++)
        if (rp_size[k] < restart_size) restart_size = rp_size[k];
      step_len = first_size;
      return enter(kDescent);
    }
  }
  if (!rp_started) {
    rp_started = true;
    for (int i = 0; i < n_par; i++) {
      rp_live[i] = 1;
      rp_move[i] = -rp_size[i] * sgn(REAL(1));
      next[i] = next[i] + rp_move[i];
    }
    return next;
  }
  if (have < 2) PPH_ERROR("RPROP called with no gradient");
  const REAL* g_prev = slopes.newest(1);
  const REAL* g_prev2 = slopes.newest(have >= 3 ? 2 : have - 1);      // only read when three gradients exist
  for (int i = 0; i < n_par; i++) {
    const int agreement = (int)((REAL)rp_live[i] * sgn(g[i] * g_prev[i]));
    const bool uncapped = (i == n_par - 1);
    if (agreement < 0) {
      rp_live[i] = 0;
      continue;
    }
    if (agreement > 0) {
      rp_size[i] *= 1.2;
      if (rp_size[i] > size_ceiling && !uncapped) rp_size[i] = size_ceiling;
    } else if (have >= 3) {
      const REAL curvature = (g[i] - g_prev2[i]) / rp_move[i];
      if (curvature > 0) rp_size[i] = std::fabs(g[i]) / curvature;
      if (rp_size[i] < size_floor) rp_size[i] = size_floor;
      if (rp_size[i] > size_ceiling && !uncapped) rp_size[i] = size_ceiling;
      rp_live[i] = 1;
    } else {
      continue;
    }
    rp_move[i] = -sgn(g[i]) * rp_size[i];
    next[i] = next[i] + rp_move[i];
  }
  return next;
}

// ---- steepest descent with an adaptive step (Optimizer.cpp:400-449) ----------------------------------------------------------
std::vector<REAL> Optimizer::step_descent()
{
  std::vector<REAL> next(points.newest(0), points.newest(0) + n_par);
  const REAL* g = slopes.newest(0);
  const REAL len = length(g, n_par);
  dir.assign(g, g + n_par);
  for (int i = 0; i < n_par; i++) dir[i] /= len;
  dir_set = true;
  const int have = errors.size();
  if (have == 1) {
    step_len = 5e-5;
    for (int i = 0; i < n_par; i++) next[i] -= step_len * dir[i];
    return next;
  }
  step_len *= (*errors.newest(0) <= *errors.newest(1)) ? 1.4 : 0.7;
  if (step_len > size_ceiling) step_len = size_ceiling;
  else if (step_len < size_floor) step_len = size_floor;
  if (may_switch && epoch % 50 == 0) {
    // default schedule: relative drift of |g| over the kept trajectory decides between L-BFGS (slowly falling) and Rprop (rising)
    REAL drift = 0.0;
    REAL total = length(slopes.oldest(0), n_par);
    for (int i = 1; i < slopes.size(); i++) {
      total += length(slopes.oldest(i), n_par);
      drift += length(slopes.oldest(i), n_par) - length(slopes.oldest(i - 1), n_par);
    }
    drift /= total;
    if (drift > -F.debug && drift < 0) return enter(kLbfgs);
    if (drift > F.debug) return enter(kRprop);
  }
  const REAL* x_now = points.newest(0);
  const REAL* x_before = points.newest(1);
  for (int i = 0; i < n_par; i++) next[i] = next[i] - step_len * (dir[i] + F.sd_momentum * (x_now[i] - x_before[i]));
  return next;
}

// switching between the methods of the default schedule (Optimizer.h:241-268)
std::vector<REAL> Optimizer::enter(Method m)
{
  static const char* banner[] = {"----- Switching to rprop -----", "----- Switching to SD -----", "", "----- Switching to BFGS -----"};
  if (mpi->io_node() && do_print) std::cout << banner[m] << std::endl;
  method = m;
  dir.clear();
  dir_set = false;
  rp_started = false;
  if (m == kRprop) { line_done = true; rp_size.assign(n_par, restart_size); }
  if (m == kDescent) line_done = true;
  return dispatch();
}

std::vector<REAL> Optimizer::dispatch()
{
  switch (method) {
    case kRprop: return step_rprop();
    case kDescent: return step_descent();
    case kDescentLine: return step_descent_line();
    default: return step_lbfgs();
  }
}

// the point at distance t along -dir from the line's origin
std::vector<REAL> Optimizer::along_line(REAL t) const
{
  std::vector<REAL> p = line_origin;
  for (int i = 0; i < n_par; i++) p[i] -= t * dir[i];
  return p;
}

void Optimizer::start_line_at_current(const REAL* direction_source)
{
  unit_vector(direction_source, n_par, dir);
  dir_set = true;
  line_origin.assign(points.newest(0), points.newest(0) + n_par);
  line_origin_slope.assign(slopes.newest(0), slopes.newest(0) + n_par);
}

// ---- steepest descent, one line minimisation per direction (Optimizer.cpp:457-491) ----------------------------------------------
std::vector<REAL> Optimizer::step_descent_line()
{
  if (!dir_set || line_done) start_line_at_current(slopes.newest(0));
  return along_line(line_search(F.line_min_epsilon));
}

// ---- L-BFGS, two-loop recursion over at most lbfgs_memory pairs (Optimizer.cpp:607-696) ---------------------------------------------
std::vector<REAL> Optimizer::step_lbfgs()
{
  if (!dir_set) {
    start_line_at_current(slopes.newest(0));
    return along_line(line_search(F.line_min_epsilon));
  }
  if (!line_done) return along_line(line_search(F.line_min_epsilon));

  const REAL* x_now = points.newest(0);
  const REAL* g_now = slopes.newest(0);
  if (epoch % (8 * lbfgs_memory) == 0) pairs.clear();
  Pair fresh;
  fresh.s.resize(n_par);
  fresh.y.resize(n_par);
  for (int i = 0; i < n_par; i++) {
    fresh.s[i] = x_now[i] - line_origin[i];
    fresh.y[i] = g_now[i] - line_origin_slope[i];
  }
  fresh.rho = 1 / inner(fresh.y, fresh.s);
  pairs.push_back(fresh);
  if ((int)pairs.size() > lbfgs_memory) pairs.erase(pairs.begin());
  const int m = (int)pairs.size();
  std::vector<REAL> coeff(m + 1, 0.0);
  std::vector<REAL> d(g_now, g_now + n_par);
  for (int k = m - 1; k >= 0; k--) {
    coeff[k] = pairs[k].rho * inner(pairs[k].s, d);
    for (int j = 0; j < n_par; j++) d[j] -= coeff[k] * pairs[k].y[j];
  }
  const Pair& last = pairs.back();
  const REAL scale = inner(last.y, last.s) / inner(last.y, last.y);
  for (int j = 0; j < n_par; j++) d[j] *= scale;
  for (int k = 0; k < m; k++) {
    const REAL back = pairs[k].rho * inner(pairs[k].y, d);
    for (int j = 0; j < n_par; j++) d[j] += pairs[k].s[j] * (coeff[k] - back);
  }
  start_line_at_current(d.data());
  return along_line(line_search(F.line_min_epsilon));
}

// ---- line minimisation on the directional slope (Optimizer.cpp:709-869) ----------------------------------------------------------
// Driven one gradient evaluation per call; returns the distance along -dir from line_origin at which the next gradient
// is wanted.  Fresh: take the trial step first_size.  Expanding: one reference point, keep doubling (or take a secant
// step once the slope flattens) while the error falls.  Bracketed: two points whose slopes differ in sign, [0] the end
// with the larger |slope|; inverse quadratic interpolation through the three most recent slopes, secant when two of
// them coincide, bisection whenever the interpolant misbehaves.
REAL Optimizer::line_search(REAL tolerance)
{
  const REAL eps = std::numeric_limits<REAL>::epsilon();
  std::vector<REAL> u;
  unit_vector(dir.data(), n_par, u);
  const REAL slope = inner(u.data(), slopes.newest(0), n_par);
  const REAL E_now = *errors.newest(0);

  if (line_done || ls.npts == 0) {
    line_done = false;
    ls.phase = LineSearch::kExpanding;
    ls.npts = 1;
    ls.a[0] = 0.0; ls.s[0] = slope;
    step_len = first_size;
    ls.ok_E = E_now; ls.ok_a = step_len; ls.ok_s = slope;
    return step_len;
  }

  const REAL a_front = ls.a[0], a_back = ls.a[ls.npts - 1];
  const bool collapsed = std::abs(a_front - a_back) < 20 * eps;
  if (std::abs(slope) < tolerance || (ls.npts == 2 && collapsed)) {
    line_done = true;
    if (collapsed && std::abs(slope) > tolerance)
      std::cout << "<<WARNING>> In BFGS: Line optimization could not reach the requested precision" << std::endl;
    return step_len;
  }

  if (ls.npts == 1) {
    if (sgn(slope * ls.s[0]) > 0) {
      // same sign as the reference point: not bracketed yet
      if (E_now <= ls.ok_E) {
        const REAL a_ref = ls.a[0], s_ref = ls.s[0];
        ls.ok_s = slope; ls.ok_a = step_len; ls.ok_E = E_now;
        const REAL a_here = step_len;
        if (std::abs(slope) >= std::abs(s_ref)) step_len *= 2.0;
        else step_len = step_len - slope * (a_here - a_ref) / (slope - s_ref);
        ls.a[0] = a_here; ls.s[0] = slope;
      } else if (step_len < 10 * first_size) {
        // the error rose within a few trial steps: give this direction up and restart from the plain gradient
        line_done = true;
        pairs.clear();
        start_line_at_current(slopes.newest(0));
      } else {
        // overshot: back to the last accepted point, creep forward by the trial step
        step_len = ls.ok_a + sgn(ls.ok_a) * first_size;
        ls.a[0] = ls.ok_a; ls.s[0] = ls.ok_s;
      }
      return step_len;
    }
    // the slope changed sign: bracketed.  [1] takes the point with the smaller |slope|
    if (std::abs(slope) <= std::abs(ls.s[0])) { ls.a[1] = step_len; ls.s[1] = slope; }
    else { ls.a[1] = ls.a[0]; ls.s[1] = ls.s[0]; ls.a[0] = step_len; ls.s[0] = slope; }
    ls.npts = 2;
    ls.phase = LineSearch::kBracketed;
    ls.c_a = ls.a[0]; ls.c_s = ls.s[0];
    ls.bisected = true;
  } else {
    // replace the end whose slope has the sign of the new one, keep the larger |slope| in [0]
    if (sgn(slope * ls.s[0]) < 0) { ls.s[1] = slope; ls.a[1] = step_len; }
    else { ls.s[0] = slope; ls.a[0] = step_len; }
    if (std::abs(ls.s[0]) < std::abs(ls.s[1])) { std::swap(ls.s[0], ls.s[1]); std::swap(ls.a[0], ls.a[1]); }
  }

  const REAL a0 = ls.a[0], a1 = ls.a[1], s0 = ls.s[0], s1 = ls.s[1];
  REAL t;
  if (ls.c_s != s0 && ls.c_s != s1) {
    t = a0 * s1 * ls.c_s / ((s0 - s1) * (s0 - ls.c_s))
      + a1 * s0 * ls.c_s / ((s1 - s0) * (s1 - ls.c_s))
      + ls.c_a * s0 * s1 / ((ls.c_s - s0) * (ls.c_s - s1));
  } else {
    t = (a0 * s1 - a1 * s0) / (s1 - s0);
  }
  const REAL move = std::abs(t - a1), prev_move = std::abs(a1 - ls.c_a), prev2_move = std::abs(ls.c_a - ls.d_a);
  const bool outside = (4 * t > (3 * a0 + a1) && t > a1) || (4 * t < (3 * a0 + a1) && t < a1);
  const REAL reference_move = ls.bisected ? prev_move : prev2_move;
  if (outside || 2 * move >= reference_move || reference_move < ls.resolution) {
    t = 0.5 * (a0 + a1);
    ls.bisected = true;
  } else {
    ls.bisected = false;
  }
  ls.d_a = ls.c_a;
  ls.c_a = a1;
  ls.c_s = s1;
  step_len = t;
  return t;
}

// ---- one call per epoch (Optimizer.cpp:884-1064) ----------------------------------------------------------------------------
void Optimizer::update_network(REAL Error, const std::vector<REAL>& grad)
{
  // Error and grad arrive already summed over ranks
  std::copy(grad.begin(), grad.end(), slopes.append());
  *errors.append() = Error / n_structures;
  REAL shown_E = *errors.newest(0);                     // the table shows the error without the L2 term
  if (F.regularization) {
    REAL* g = slopes.newest_mut();
    REAL* e = errors.newest_mut();
    const REAL* x = points.newest(0);
    for (int i = 0; i < n_par - 1; i++) {               // the last parameter (an output bias) is left out
      g[i] += 2 * F.regularization * x[i] / n_par;
      *e += F.regularization * x[i] * x[i] / n_par / n_structures;
    }
  }
  REAL shown_norm = length(slopes.newest(0), n_par) / n_par / n_structures;

  done = false;
  if (epoch >= F.Niterations) {
    done = true;
    line_done = true;
  } else if (shown_norm < F.threshold) {
    if (pt.ceiling == 0) {
      done = true;
      line_done = true;
    } else if (pt.alpha == 0) {
      // lambda_max given: instead of stopping, a slow sine of that amplitude is switched on for half a period.
      // Potential::train never reads alpha(), so on this path the scheme only decides WHEN training stops; it is kept
      // for that and for the random numbers it draws.
      if (pt.enabled) {
        if (shown_E < pt.best_E) { pt.best_E = shown_E; pt.best_norm = shown_norm; pt.best_x.assign(points.newest(0), points.newest(0) + n_par); }
        if (mpi->io_node()) std::cout << "----- converged; turning on perturbation -----" << std::endl;
        pt.floor_E = shown_E;
        pt.rate = 3.14159265358979e-04;
        pt.t_on = epoch;
      } else {
        done = true;
        line_done = true;
        if (shown_E > pt.best_E) {
          if (mpi->io_node()) std::cout << "===== resetting to previous best =====" << std::endl;
          shown_E = pt.best_E;
          shown_norm = pt.best_norm;
          std::copy(pt.best_x.begin(), pt.best_x.end(), points.append());      // the caller reads parameters()
        }
      }
    }
  }
  if (!done) {
    const std::vector<REAL> next = dispatch();
    std::copy(next.begin(), next.end(), points.append());
  }
  if (points.size() > kKeep) {
    points.drop_oldest();
    slopes.drop_oldest();
    errors.drop_oldest();
  }
  current.assign(points.newest(0), points.newest(0) + n_par);

  if (line_done) {
    if (do_print && mpi->io_node() && (epoch % F.Nprint == 0 || done)) {
      if (!report_validation) printf("%-9d %#13.6e %#13.6e\n", epoch, std::sqrt(shown_E), shown_norm);
      else printf("%-12d %#13.6e    %#13.6e    %#13.6e\n", epoch, std::sqrt(shown_E), std::sqrt(val_sse / (REAL)val_count), shown_norm);
      fflush(stdout);
    }
    epoch++;
    want_checkpoint = ((epoch - 1) % F.Ncheckpoint == 0);
    if (F.Tbackup) want_backup = ((epoch - 1) % F.Nbackup == 0);
    if (epoch == 1) { want_checkpoint = false; want_backup = false; }
  }
  if (do_print && pt.rate) perturbation_tick(shown_E);      // inside the reference's do_print branch (Optimizer.cpp:1008-1043)
}

// the slow sine of the perturbation scheme: amplitude, a noisy phase rate, and the decision to go on or to finish
void Optimizer::perturbation_tick(REAL shown_E)
{
  if (shown_E < 0.985 * pt.floor_E && pt.phase == 0) {
    if (mpi->io_node()) std::cout << "----- transition detected -----" << std::endl;
    pt.amplitude_now = pt.alpha;
    pt.t_on = epoch;
    pt.cycles = 0;
    pt.phase = 1.57079632679490;
    pt.rate *= 2;
  }
  pt.alpha = pt.amplitude_now * std::sin(pt.rate * (epoch - pt.t_on) + pt.phase);
  pt.dphi = 0.5 * gaussian_from_rand(1, 0.15) * 0.0314159 / 5;
  if (pt.alpha > 0) return;
  if (pt.amplitude_now == pt.ceiling) {
    if (mpi->io_node()) std::cout << "----- No transition found; completing optimization -----" << std::endl;
    pt.cycles++;
    if (pt.cycles == F.Ncycles) pt.enabled = false;
  } else {
    if (mpi->io_node()) std::cout << "----- continuing optimization -----" << std::endl;
    pt.phase = 0;
    pt.amplitude_now = pt.ceiling;
  }
  pt.alpha = 0;
  pt.rate = 0;
}

}  // namespace pph
