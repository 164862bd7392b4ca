// Optimizer of the host layer: Optimizer::update_network and the Rprop / steepest-descent steps
// (reference: Optimizer.cpp:193-289, 400-449, 521-593, 884-1064).  It runs on the host on the already
// reduced gradient (P <= ~50k doubles): every rank holds the identical allreduced vector, so every rank takes the
// identical step and the reference's Bcast of the new parameters (Optimizer.cpp:957) is not needed.
#include <cmath>
#include <cstdio>
#include <limits>

#include "pph.h"

namespace pph {

template <typename T>
static inline int sign(T v) { return (T(0) < v) - (v < T(0)); }

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
  else PPH_ERROR("Training algorithm \"" + a + "\" is not implemented in this build (default, rprop, steepest_descent are)");
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
        REAL D2 = (g[i] - gradient[N - 3][i]) / rprop_step[i];
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
    // the reference would hand over to L-BFGS when dg in (-debug, 0); that optimizer is not part of this build,
    // so steepest descent simply continues in that case
    if (dg > debug) return switch_to_rprop();
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

std::vector<REAL> Optimizer::next_x() { return algo == RPROP ? rprop() : steepest_descent(); }

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
    my_is_converged = true;       // lambda_max = 0: no perturbation phase (the "global optimizer" is not in this build)
    line_min = true;
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
}

}  // namespace pph
