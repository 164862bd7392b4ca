// Optimizer of the host layer: Optimizer::update_network and the Rprop / steepest-descent steps
// (reference: Optimizer.cpp:193-289, 400-449, 521-593, 884-1064).  It runs on the host on the already
// reduced gradient (P <= ~50k doubles): every rank holds the identical allreduced vector, so every rank takes the
// identical step and the reference's Bcast of the new parameters (Optimizer.cpp:957) is not needed.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <limits>

#include "pph.h"

namespace pph {

template <typename T>
static inline int sign(T v) { return (T(0) < v) - (v < T(0)); }

// RAND::Normal (Common.h:72-75): libc rand(), first draw under the logarithm, second in the cosine
static REAL rand_normal_opt(REAL mean, REAL sd)
{
  const REAL u1 = (REAL)rand() / RAND_MAX;
  const REAL u2 = (REAL)rand() / RAND_MAX;
  return mean + sd * std::sqrt(-2.0 * std::log(u1)) * std::cos(2.0 * 3.141592653589793 * u2);
}

static REAL dot(const std::vector<REAL>& a, const std::vector<REAL>& b)
{
  REAL r = 0.0;
  for (size_t i = 0; i < a.size(); i++) r += a[i] * b[i];
  return r;
}

Optimizer::Optimizer(Parallel* mpi_in, const FunctionalParams& params, int Nsystems) : mpi(mpi_in), F(params)
{
  do_print = true;
  early_stop = false;
  my_Nsystems = Nsystems;
  my_max_steps = params.Niterations;
  my_Ncheckpoint = params.Ncheckpoint;
  my_Tbackup = params.Tbackup;
  my_Nbackup = params.Nbackup;
  my_threshold = params.threshold;
  debug = params.debug;
  // Optimizer::set_training_algorithm (Optimizer.cpp:139-166)
  const std::string& a = params.training_algorithm;
  if (a == "default") { algo = RPROP; allow_training_switch = true; }
  else if (a == "rprop") { algo = RPROP; allow_training_switch = false; }
  else if (a == "steepest_descent") { algo = SD; allow_training_switch = false; }
  else if (a == "steepest_descent_with_line_opt") { algo = SD_LINE; allow_training_switch = false; }
  else if (a == "bfgs") { algo = BFGS; allow_training_switch = false; }
  else PPH_ERROR("Training algorithm \"" + a + "\" is not implemented");
  iteration_counter = 1;
  my_is_converged = my_is_checkpoint = my_is_backup = false;
  line_min = true;
  SSE_val = 0;
  nval = 0;
  my_Nparameters = 0;
  lambda0 = 1e-4;
  lambda = lambda0;
  alpha1 = lambda0;
  rprop_weights_max = 0.5;
  rprop_weights_min = 1000.0 * std::numeric_limits<REAL>::epsilon();
  my_alpha = 0.0;
  my_dphi_dt = 0.0;
  global_optimizer = true;
  alpha_max = (int)F.lambda_max;      // the reference's accessor returns int (Functional_params.h:198-201): values below 1 switch the scheme off
  current_alpha_max = alpha_max;
  phase_shift = 0.0;
  dalpha_dt = 0;
  cycles = 0;
  t0 = 0;
  E_best = 9e9;
  g_best = 0.0;
  E_min = 0.0;
  BFGS_Nmax = 77;
  convergence = 100 * std::numeric_limits<REAL>::epsilon();
  line_last_E = line_last_alpha = line_last_g = brent_c = brent_g_c = brent_d = 0.0;
  brent_flag = false;
}

void Optimizer::init(const std::vector<REAL>& x0)
{
  // Optimizer::internal_init (Optimizer.cpp:242-289)
  if (mpi->io_node() && do_print) {
    if (early_stop) {
      std::cout << "Iteration    Ɛ_train = Σ(E²)    Ɛ_val = Σ(E²)       |∇Ɛ|" << std::endl;
      std::cout << "-------------------------------------------------------------" << std::endl;
    } else {
      std::cout << "Iteration    Ɛ = Σ(E²)       |∇Ɛ|" << std::endl;
      std::cout << "-------------------------------------" << std::endl;
    }
  }
  x.clear(); gradient.clear(); Errors.clear();
  x.push_back(x0);
  my_Nparameters = (int)x0.size();
  rprop_weights.assign(my_Nparameters, lambda0);
  rprop_step.assign(my_Nparameters, 0);
  search_direction.clear();
}

std::vector<REAL> Optimizer::rprop()
{
  // Optimizer.cpp:521-593
  line_min = true;
  const int N = (int)gradient.size();
  std::vector<REAL> x1 = x.back();
  if (allow_training_switch && !(iteration_counter % 67)) {
    if (gradient.size() > 1 && dot(gradient.back(), gradient.back()) < 1e11 && search_direction.size() > 0) {
      const std::vector<REAL>& g = gradient.back();
      const std::vector<REAL>& gp = gradient[gradient.size() - 2];
      for (size_t i = 0; i < g.size(); i++) {
        if ((g[i] - gp[i]) / rprop_step[i] < 0) break;
        if (i == g.size() - 1) {
          for (size_t k = 0; k < g.size(); k++)
            if (rprop_weights[k] < lambda) lambda = rprop_weights[k];
          alpha1 = lambda0;
          return switch_to_sd();
        }
      }
    }
  }
  const std::vector<REAL>& g = gradient.back();
  const int P = (int)g.size();
  if (search_direction.size() == 0) {
    search_direction.assign(P, 1);
    for (int i = 0; i < P; i++) {
      rprop_step[i] = -rprop_weights[i] * sign(search_direction[i]);
      x1[i] = x1[i] + rprop_step[i];
    }
  } else if (N > 1) {
    const std::vector<REAL>& g1 = gradient[N - 2];
    const std::vector<REAL>& g2 = gradient[N >= 3 ? N - 3 : 0];      // only read when N >= 3
    for (int i = 0; i < P; i++) {
      int sign_change = (int)(search_direction[i] * sign(g[i] * g1[i]));
      if (sign_change < 0) {
        search_direction[i] = 0;
      } else if (sign_change > 0) {
        rprop_weights[i] *= 1.2;
        if (rprop_weights[i] > rprop_weights_max && i != P - 1) rprop_weights[i] = rprop_weights_max;
        rprop_step[i] = -sign(g[i]) * rprop_weights[i];
        x1[i] = x1[i] + rprop_step[i];
      } else if (sign_change == 0 && N >= 3) {
        REAL D2 = (g[i] - g2[i]) / rprop_step[i];
        if (D2 > 0) rprop_weights[i] = std::fabs(g[i]) / D2;
        if (rprop_weights[i] < rprop_weights_min) rprop_weights[i] = rprop_weights_min;
        if (rprop_weights[i] > rprop_weights_max && i != P - 1) rprop_weights[i] = rprop_weights_max;
        rprop_step[i] = -sign(g[i]) * rprop_weights[i];
        x1[i] = x1[i] + rprop_step[i];
        search_direction[i] = 1;
      }
    }
  } else {
    PPH_ERROR("RPROP called with no gradient");
  }
  return x1;
}

std::vector<REAL> Optimizer::steepest_descent()
{
  // Optimizer.cpp:400-449
  std::vector<REAL> x1 = x.back();
  const std::vector<REAL>& g = gradient.back();
  REAL norm = std::sqrt(dot(g, g));
  search_direction = g;
  for (size_t i = 0; i < g.size(); i++) search_direction[i] /= norm;
  const int N = (int)Errors.size();
  if (N == 1) {
    alpha1 = 5e-5;
    for (size_t i = 0; i < x1.size(); i++) x1[i] -= alpha1 * search_direction[i];
    return x1;
  } else if (Errors[N - 1] <= Errors[N - 2]) {
    alpha1 *= 1.4;
  } else {
    alpha1 *= 0.7;
  }
  if (alpha1 > rprop_weights_max) alpha1 = rprop_weights_max;
  else if (alpha1 < rprop_weights_min) alpha1 = rprop_weights_min;
  if (allow_training_switch && !(iteration_counter % 50)) {
    REAL dg = 0.0;
    REAL ave = std::sqrt(dot(gradient[0], gradient[0]));
    for (size_t i = 1; i < gradient.size(); i++) {
      ave += std::sqrt(dot(gradient[i], gradient[i]));
      dg += std::sqrt(dot(gradient[i], gradient[i])) - std::sqrt(dot(gradient[i - 1], gradient[i - 1]));
    }
    dg /= ave;
    if (dg > -debug && dg < 0) return switch_to_BFGS();
    else if (dg > debug) return switch_to_rprop();
  }
  for (size_t i = 0; i < x1.size(); i++)
    x1[i] = x1[i] - alpha1 * (search_direction[i] + F.sd_momentum * (x[N - 1][i] - x[N - 2][i]));
  return x1;
}

std::vector<REAL> Optimizer::switch_to_sd()
{
  if (mpi->io_node() && do_print) std::cout << "----- Switching to SD -----" << std::endl;      // Optimizer.h:253-260
  line_min = true;
  search_direction.clear();
  algo = SD;
  return steepest_descent();
}

std::vector<REAL> Optimizer::switch_to_rprop()
{
  search_direction.clear();                                                                    // Optimizer.h:241-251
  if (mpi->io_node() && do_print) std::cout << "----- Switching to rprop -----" << std::endl;
  line_min = true;
  rprop_weights.assign(my_Nparameters, lambda);
  algo = RPROP;
  return rprop();
}

std::vector<REAL> Optimizer::switch_to_BFGS()
{
  if (mpi->io_node() && do_print) std::cout << "----- Switching to BFGS -----" << std::endl;     // Optimizer.h:262-268
  search_direction.clear();
  algo = BFGS;
  return LBFGS();
}

std::vector<REAL> Optimizer::next_x()
{
  switch (algo) {
    case RPROP: return rprop();
    case SD: return steepest_descent();
    case SD_LINE: return steepest_descent_with_line_opt();
    default: return LBFGS();
  }
}

static std::vector<REAL> normalized(const std::vector<REAL>& v)
{
  std::vector<REAL> out = v;
  REAL norm = std::sqrt(dot(v, v));
  for (size_t i = 0; i < v.size(); i++) out[i] /= norm;
  return out;
}

std::vector<REAL> Optimizer::steepest_descent_with_line_opt()
{
  // Optimizer.cpp:457-491
  if (search_direction.size() == 0 || line_min) {
    std::vector<REAL> x1 = x.back();
    search_direction = normalized(gradient.back());
    BFGS_last_x = x.back();
    BFGS_last_gradient = gradient.back();
    REAL delta = brent_line_min(F.line_min_epsilon);
    for (size_t i = 0; i < x1.size(); i++) x1[i] -= delta * search_direction[i];
    return x1;
  }
  std::vector<REAL> x1 = BFGS_last_x;
  REAL delta = brent_line_min(F.line_min_epsilon);
  for (size_t i = 0; i < search_direction.size(); i++) x1[i] -= delta * search_direction[i];
  return x1;
}

std::vector<REAL> Optimizer::LBFGS()
{
  // Optimizer.cpp:607-696
  if (search_direction.size() == 0) {
    std::vector<REAL> x1 = x.back();
    search_direction = normalized(gradient.back());
    BFGS_last_x = x.back();
    BFGS_last_gradient = gradient.back();
    REAL delta = brent_line_min(F.line_min_epsilon);
    for (size_t i = 0; i < x1.size(); i++) x1[i] -= delta * search_direction[i];
    return x1;
  }
  if (!line_min) {
    std::vector<REAL> x1 = BFGS_last_x;
    REAL delta = brent_line_min(F.line_min_epsilon);
    for (size_t i = 0; i < search_direction.size(); i++) x1[i] -= delta * search_direction[i];
    return x1;
  }
  std::vector<REAL> alpha(s.size() + 1, 0.0);
  REAL beta;
  search_direction = gradient.back();
  if (!(iteration_counter % (8 * BFGS_Nmax))) { s.clear(); y.clear(); rho.clear(); }
  const size_t P = search_direction.size();
  s.push_back(std::vector<REAL>(P, 0.0));
  y.push_back(std::vector<REAL>(P, 0.0));
  for (size_t i = 0; i < P; i++) {
    s.back()[i] = x.back()[i] - BFGS_last_x[i];
    y.back()[i] = gradient.back()[i] - BFGS_last_gradient[i];
  }
  rho.push_back(1 / dot(y.back(), s.back()));
  if ((int)s.size() > BFGS_Nmax) { s.erase(s.begin()); y.erase(y.begin()); rho.pop_front(); }
  for (int i = (int)s.size() - 1; i >= 0; i--) {
    alpha.at(i) = rho[i] * dot(s[i], search_direction);
    for (size_t j = 0; j < P; j++) search_direction[j] -= alpha[i] * y[i][j];
  }
  REAL H = dot(y.back(), s.back()) / dot(y.back(), y.back());
  for (size_t j = 0; j < P; j++) search_direction[j] *= H;
  for (size_t i = 0; i < s.size(); i++) {
    beta = rho[i] * dot(y[i], search_direction);
    for (size_t j = 0; j < P; j++) search_direction[j] += s[i][j] * (alpha[i] - beta);
  }
  search_direction = normalized(search_direction);
  BFGS_last_x = x.back();
  BFGS_last_gradient = gradient.back();
  REAL delta = brent_line_min(F.line_min_epsilon);
  std::vector<REAL> x1 = BFGS_last_x;
  for (size_t i = 0; i < P; i++) x1[i] -= delta * search_direction[i];
  return x1;
}


REAL Optimizer::brent_line_min(REAL tolerance)
{
  // Optimizer.cpp:709-869.  The line search is driven one gradient evaluation per call: the bracket lives in the
  // two-slot deques line_alpha / line_g (front = the end with the larger |g| once a sign change is bracketed),
  // (brent_c, brent_g_c) is the previous best point and brent_d the one before it.  The value returned is the
  // step length along -search_direction measured from BFGS_last_x.
  const REAL eps = std::numeric_limits<REAL>::epsilon();
  const REAL g = dot(normalized(search_direction), gradient.back());
  // ---- a fresh line: take the trial step lambda0 ----
  if (line_min || line_g.empty()) {
    line_min = false;
    line_g.assign(1, g);
    line_alpha.assign(1, 0.0);
    alpha1 = lambda0;
    line_last_E = Errors.back();
    line_last_alpha = alpha1;
    line_last_g = g;
    return alpha1;
  }

  // ---- finished: slope small enough, or the bracket has collapsed ----
  const bool collapsed = std::abs(line_alpha.front() - line_alpha.back()) < 20 * eps;
  if (std::abs(g) < tolerance || (line_g.size() == 2 && collapsed)) {
    line_min = true;
    if (collapsed && std::abs(g) > tolerance)
      std::cout << "<<WARNING>> In BFGS: Line optimization could not reach the requested precision" << std::endl;
    return alpha1;
  }

  if (line_g.size() == 1) {
    if (sign(g * line_g.back()) <= 0) {
      // the slope changed sign: the minimum is bracketed; keep the smaller |g| at the back
      if (std::abs(g) <= std::abs(line_g.back())) { line_g.push_back(g); line_alpha.push_back(alpha1); }
      else { line_g.push_front(g); line_alpha.push_front(alpha1); }
      brent_c = line_alpha.front();
      brent_g_c = line_g.front();
      brent_flag = true;
    } else {
      // still going downhill along the line: expand (double, or secant when the slope is flattening)
      if (Errors.back() <= line_last_E) {
        line_alpha.push_back(alpha1);
        line_g.push_back(g);
        line_last_g = g;
        line_last_alpha = alpha1;
        line_last_E = Errors.back();
        if (std::abs(g) >= std::abs(line_g.front())) alpha1 *= 2.0;
        else alpha1 = alpha1 - g * (line_alpha.back() - line_alpha.front()) / (line_g.back() - line_g.front());
        line_alpha.pop_front();
        line_g.pop_front();
      } else if (alpha1 < 10 * lambda0) {
        // the error rose within a few trial steps: give the direction up and restart from the plain gradient
        line_min = true;
        s.clear();
        y.clear();
        rho.clear();
        search_direction = normalized(gradient.back());
        BFGS_last_x = x.back();
        BFGS_last_gradient = gradient.back();
      } else {
        // overshot: go back to the last accepted point and creep forward by lambda0
        alpha1 = line_last_alpha + sign(line_last_alpha) * lambda0;
        line_alpha.front() = line_last_alpha;
        line_g.front() = line_last_g;
      }
      return alpha1;
    }
  } else if (line_g.size() == 2) {
    // replace the bracket end whose slope has the sign of the new one
    if (sign(g * line_g.front()) < 0) { line_g.back() = g; line_alpha.back() = alpha1; }
    else { line_g.front() = g; line_alpha.front() = alpha1; }
    if (std::abs(line_g.front()) < std::abs(line_g.back())) {
      std::swap(line_g.front(), line_g.back());
      std::swap(line_alpha.front(), line_alpha.back());
    }
  } else {
    PPH_ERROR("Something went wrong");
  }
  if (line_g.size() > 2 || line_alpha.size() > 2) PPH_ERROR("Incorrect BFGS");

  // ---- next trial: inverse quadratic interpolation through three slopes, else secant; bisection as the safeguard
  const REAL a0 = line_alpha.front(), a1 = line_alpha.back(), g0 = line_g.front(), g1 = line_g.back();
  REAL trial;
  if (brent_g_c != g0 && brent_g_c != g1) {
    trial = a0 * g1 * brent_g_c / ((g0 - g1) * (g0 - brent_g_c))
          + a1 * g0 * brent_g_c / ((g1 - g0) * (g1 - brent_g_c))
          + brent_c * g0 * g1 / ((brent_g_c - g0) * (brent_g_c - g1));
  } else {
    trial = (a0 * g1 - a1 * g0) / (g1 - g0);
  }
  const REAL da = std::abs(trial - a1), db = std::abs(a1 - brent_c), dc = std::abs(brent_c - brent_d);
  const bool outside = (4 * trial > (3 * a0 + a1) && trial > a1) || (4 * trial < (3 * a0 + a1) && trial < a1);
  if (outside || (brent_flag && 2 * da >= db) || (!brent_flag && 2 * da >= dc) ||
      (brent_flag && db < convergence) || (!brent_flag && dc < convergence)) {
    trial = 0.5 * (a0 + a1);
    brent_flag = true;
  } else {
    brent_flag = false;
  }
  brent_d = brent_c;
  brent_c = a1;
  brent_g_c = g1;
  alpha1 = trial;
  return trial;
}

void Optimizer::update_network(REAL Error, const std::vector<REAL>& grad)
{
  // Optimizer.cpp:884-1064.  Error and grad arrive already summed over ranks.
  Errors.push_back(Error / my_Nsystems);
  gradient.push_back(grad);
  REAL SSE = Errors.back();
  if (F.regularization) {
    std::vector<REAL>& g = gradient.back();
    const std::vector<REAL>& xx = x.back();
    for (int i = 0; i < my_Nparameters - 1; i++) {
      g[i] += 2 * F.regularization * xx[i] / my_Nparameters;
      Errors.back() += F.regularization * xx[i] * xx[i] / my_Nparameters / my_Nsystems;
    }
  }
  REAL norm = std::sqrt(dot(gradient.back(), gradient.back())) / my_Nparameters / my_Nsystems;
  my_is_converged = false;
  if (iteration_counter >= my_max_steps) {
    my_is_converged = true;
    line_min = true;
  } else if (norm < my_threshold) {
    if (alpha_max != 0) {
      // lambda_max given: instead of stopping, a slow sine of amplitude lambda_max is switched on for half a period
      // (Optimizer.cpp:915-938).  Potential::train never reads alpha(), so on this path the scheme only decides WHEN
      // training stops; it is reproduced for that and for the random numbers it draws.
      if (my_alpha == 0) {
        if (global_optimizer) {
          if (SSE < E_best) { E_best = SSE; g_best = norm; x_best = x.back(); }
          if (mpi->io_node()) std::cout << "----- converged; turning on perturbation -----" << std::endl;
          E_min = SSE;
          dalpha_dt = 3.14159265358979e-04;
          t0 = iteration_counter;
        } else {
          my_is_converged = true;
          line_min = true;
          if (SSE > E_best) {
            if (mpi->io_node()) std::cout << "===== resetting to previous best =====" << std::endl;
            SSE = E_best;
            norm = g_best;
            x.push_back(x_best);      // set_parameters(x_best): the caller reads parameters()
          }
        }
      }
    } else {
      my_is_converged = true;
      line_min = true;
    }
  }
  if (!my_is_converged) x.push_back(next_x());
  if (x.size() > 20) {
    x.pop_front();
    gradient.pop_front();
    Errors.pop_front();
  }
  REAL sse_val = 0;
  if (early_stop) sse_val = std::sqrt(SSE_val / (REAL)nval);
  if (line_min) {
    if (do_print && mpi->io_node()) {
      if (!(iteration_counter % F.Nprint) || my_is_converged) {
        if (!early_stop) printf("%-9d %#13.6e %#13.6e\n", iteration_counter, std::sqrt(SSE), norm);
        else printf("%-12d %#13.6e    %#13.6e    %#13.6e\n", iteration_counter, std::sqrt(SSE), sse_val, norm);
        fflush(stdout);
      }
    }
    iteration_counter++;
    my_is_checkpoint = ((iteration_counter - 1) % my_Ncheckpoint ? false : true);
    if (my_Tbackup) my_is_backup = ((iteration_counter - 1) % my_Nbackup ? false : true);
    if (iteration_counter == 1) { my_is_checkpoint = false; my_is_backup = false; }
  }
  if (do_print && dalpha_dt) {      // Optimizer.cpp:1008-1043 (inside the reference's do_print branch)
    if (SSE < 0.985 * E_min && phase_shift == 0) {
      if (mpi->io_node()) std::cout << "----- transition detected -----" << std::endl;
      current_alpha_max = my_alpha;
      t0 = iteration_counter;
      cycles = 0;
      phase_shift = 1.57079632679490;
      dalpha_dt *= 2;
    }
    my_alpha = current_alpha_max * std::sin(dalpha_dt * (iteration_counter - t0) + phase_shift);
    my_dphi_dt = 0.5 * rand_normal_opt(1, 0.15) * 0.0314159 / 5;
    if (my_alpha <= 0) {
      if (current_alpha_max == alpha_max) {
        if (mpi->io_node()) std::cout << "----- No transition found; completing optimization -----" << std::endl;
        cycles++;
        if (cycles == F.Ncycles) global_optimizer = false;
        my_alpha = 0;
        dalpha_dt = 0;
      } else {
        if (mpi->io_node()) std::cout << "----- continuing optimization -----" << std::endl;
        phase_shift = 0;
        my_alpha = 0;
        current_alpha_max = alpha_max;
        dalpha_dt = 0;
      }
    }
  }
}

}  // namespace pph
