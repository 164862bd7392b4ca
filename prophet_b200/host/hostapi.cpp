// extern "C" test hooks into the host layer (used by tests/test_host_layer.py through ctypes).
#include <cstring>
#include <ctime>
#include <fstream>
#include <sstream>

#include "pph.h"

#include "pair_nn.h"

using namespace pph;

extern "C" {

// ---- Rprop driven step by step with caller-supplied gradients -------------------------------------------------
struct pph_opt { Parallel mpi; FunctionalParams F; Optimizer* opt; int nsys; };

// line_min_epsilon / debug / sd_momentum as the input_file keys of those names would set them
void* pph_optimizer_create2(int P, const double* x0, const char* algorithm, int Nsystems, double line_min_epsilon,
                            double debug, double sd_momentum);
void* pph_optimizer_create(int P, const double* x0, const char* algorithm, int Nsystems)
{
  FunctionalParams d;
  return pph_optimizer_create2(P, x0, algorithm, Nsystems, d.line_min_epsilon, d.debug, d.sd_momentum);
}
void* pph_optimizer_create2(int P, const double* x0, const char* algorithm, int Nsystems, double line_min_epsilon,
                            double debug, double sd_momentum)
{
  pph_opt* h = new pph_opt();
  h->F.training_algorithm = algorithm;
  h->F.line_min_epsilon = line_min_epsilon;
  h->F.debug = debug;
  h->F.sd_momentum = sd_momentum;
  h->F.Niterations = 1 << 30;
  h->nsys = Nsystems;
  h->opt = new Optimizer(&h->mpi, h->F, Nsystems);
  h->opt->do_print = false;
  h->opt->init(std::vector<REAL>(x0, x0 + P));
  return h;
}
// options of a test optimizer: lambda_max / Ncycles / threshold as the input_file keys, printing on (the perturbation
// scheme only runs while the optimizer prints, as in the reference); call before the first step
void pph_optimizer_configure(void* hv, double lambda_max, int ncycles, double threshold, int do_print, int P, const double* x0)
{
  pph_opt* h = (pph_opt*)hv;
  const int nsys = h->nsys;
  delete h->opt;
  h->F.lambda_max = lambda_max;
  h->F.Ncycles = ncycles;
  h->F.threshold = threshold;
  h->opt = new Optimizer(&h->mpi, h->F, nsys);
  h->opt->do_print = do_print != 0;
  h->opt->init(std::vector<REAL>(x0, x0 + P));
}
int pph_optimizer_converged(void* hv) { return ((pph_opt*)hv)->opt->is_converged() ? 1 : 0; }
double pph_optimizer_alpha(void* hv) { return ((pph_opt*)hv)->opt->alpha(); }
void pph_optimizer_step(void* hv, int P, double sse, const double* grad, double* x_out)
{
  pph_opt* h = (pph_opt*)hv;
  h->opt->update_network(sse, std::vector<REAL>(grad, grad + P));
  const std::vector<REAL>& x = h->opt->parameters();
  memcpy(x_out, x.data(), sizeof(double) * P);
}
// The body of Potential::train (pph_potential.cpp, Potential.cpp:626-672) for `steps` iterations on an existing context:
// pp_epoch with host buffers (parameters in, [gradient, SSE, vSSE, nval] out), then Optimizer::update_network.  Returns
// the wall time of the loop in seconds; x_inout holds the parameters before and after.
int pph_train_loop(void* pp_ctx, void* hv, int P, double* x_inout, double* grad_buf, int steps, double* elapsed_s, double* last_sse)
{
  pph_opt* h = (pph_opt*)hv;
  pp_context* ctx = (pp_context*)pp_ctx;
  std::vector<REAL> grad(P);
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  double sse = 0, vsse = 0;
  long long nval = 0;
  double t_opt = 0.0;
  for (int it = 0; it < steps; it++) {
    if (pp_epoch(ctx, x_inout, grad_buf, &sse, &vsse, &nval)) return 1;
    struct timespec a, b;
    clock_gettime(CLOCK_MONOTONIC, &a);
    h->opt->set_val_sse(vsse, nval);
    grad.assign(grad_buf, grad_buf + P);
    h->opt->update_network(sse, grad);
    const std::vector<REAL>& x = h->opt->parameters();
    memcpy(x_inout, x.data(), sizeof(double) * P);
    clock_gettime(CLOCK_MONOTONIC, &b);
    t_opt += (b.tv_sec - a.tv_sec) * 1e6 + 1e-3 * (b.tv_nsec - a.tv_nsec);
  }
  if (getenv("PP_EPOCH_TRACE") && steps > 0) {
    clock_gettime(CLOCK_MONOTONIC, &t1);
    fprintf(stderr, "pph_train_loop: optimizer step %.1f us, whole step %.1f us (mean of %d)\n", t_opt / steps,
            ((t1.tv_sec - t0.tv_sec) * 1e6 + 1e-3 * (t1.tv_nsec - t0.tv_nsec)) / steps, steps);
  }
  clock_gettime(CLOCK_MONOTONIC, &t1);
  if (elapsed_s) *elapsed_s = (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
  if (last_sse) *last_sse = sse;
  return 0;
}
// in-place variant for pinned buffers: x is both the current parameters' destination and not read
void pph_optimizer_step_ptr(void* hv, int P, double sse, const double* grad, double* x_inout)
{
  pph_optimizer_step(hv, P, sse, grad, x_inout);
}
void pph_optimizer_destroy(void* hv)
{
  pph_opt* h = (pph_opt*)hv;
  delete h->opt;
  delete h;
}

// ---- checkpoint: read a file, write it back ---------------------------------------------------------------------
int pph_checkpoint_roundtrip(const char* in_path, const char* out_path)
{
  std::ifstream in(in_path);
  if (!in.is_open()) return 1;
  HostNet net;
  net.params.read(in);
  net.sizes.clear();
  net.sizes.push_back(net.params.NG());
  for (size_t l = 0; l < net.params.hidden.size(); l++) net.sizes.push_back(net.params.hidden[l]);
  net.sizes.push_back(1);
  net.load(in);
  std::ofstream out(out_path);
  net.print(out);
  return 0;
}

// ---- Setup: parse an input_file (cwd must hold what it references) and report what was understood ---------------
int pph_setup_parse(const char* input_file, int rank, int nprocs, char* report, int report_len)
{
  std::ostringstream r;
  {
    std::ostringstream a, b;
    a << rank;
    b << nprocs;
    setenv("PP_RANK", a.str().c_str(), 1);
    setenv("PP_WORLD_SIZE", b.str().c_str(), 1);
  }
  Parallel mpi;
  Setup s(input_file, &mpi, true);
  const FunctionalParams& F = s.F;
  r << "Rcut " << F.Rcut << "\n";
  r << "hidden";
  for (size_t i = 0; i < F.hidden.size(); i++) r << " " << F.hidden[i];
  r << "\n";
  r << "transfer " << F.transfer_function << "\nalgorithm " << F.training_algorithm << "\n";
  r << "Niterations " << F.Niterations << "\nNprint " << F.Nprint << "\nNcheckpoint " << F.Ncheckpoint << "\n";
  r << "checkpoint_in " << F.input_file << "\ncheckpoint_out " << F.output_file << "\n";
  r << "rows " << F.G1.size() << " " << F.G2.size() << " " << F.G3.size() << " " << F.G4.size() << "\n";
  r << "Nradial " << F.Nradial << " Nangular " << F.Nangular << "\n";
  r << "systems_total " << s.Ntotal_systems << " local " << s.Nsystems() << "\n";
  for (int i = 0; i < s.Nsystems(); i++) {
    const SystemSpec& sp = s.system(i);
    r << "system";
    for (std::map<std::string, std::string>::const_iterator it = sp.files.begin(); it != sp.files.end(); ++it)
      r << " " << it->first << "=" << it->second;
    r << "\n";
  }
  std::string out = r.str();
  strncpy(report, out.c_str(), report_len - 1);
  report[report_len - 1] = 0;
  unsetenv("PP_RANK");
  unsetenv("PP_WORLD_SIZE");
  return 0;
}

// ---- readers ------------------------------------------------------------------------------------------------------
int pph_read_vasprun(const char* dir, double* cell9, int* natoms, double* pos, int* Z, double* target, int cap)
{
  StructureData sd = read_vasprun(dir);
  if (sd.natoms() > cap) return 1;
  memcpy(cell9, sd.cell, sizeof(double) * 9);
  *natoms = sd.natoms();
  memcpy(pos, sd.pos.data(), sizeof(double) * sd.pos.size());
  memcpy(Z, sd.Z.data(), sizeof(int) * sd.Z.size());
  *target = sd.target;
  return 0;
}

// Analysis::KS as Potential::validate uses it; returns D, *p_out the approximate p value
double pph_ks(const double* a, int n1, const double* b, int n2, double* p_out)
{
  REAL p = 0;
  REAL D = approximate_KS(std::vector<REAL>(a, a + n1), std::vector<REAL>(b, b + n2), &p);
  *p_out = p;
  return D;
}

int pph_auto_rows(double Rcut, int Nradial, int Nangular, double* G2_out, int* nG2, double* G4_out, int* nG4, int cap)
{
  FunctionalParams F;
  F.Rcut = Rcut;
  F.Nradial = Nradial;
  F.Nangular = Nangular;
  F.auto_rows();
  if ((int)F.G2.size() > cap || (int)F.G4.size() > cap) return 1;
  *nG2 = (int)F.G2.size();
  *nG4 = (int)F.G4.size();
  for (size_t i = 0; i < F.G2.size(); i++) memcpy(G2_out + 4 * i, F.G2[i].data(), sizeof(double) * 4);
  for (size_t i = 0; i < F.G4.size(); i++) memcpy(G4_out + 5 * i, F.G4[i].data(), sizeof(double) * 5);
  return 0;
}

// ---- pair_style nn through the LAMMPS stand-in: settings + coeff + compute on caller arrays ----------------------
int pph_pair_nn_run(double Rcut, int ntypes, const int* coeff_types, const char** coeff_files, int nlocal, int nghost,
                    double* x, int* type, const int* numneigh, const int* neigh_flat, double* f, double* eatom, double* eng_vdwl,
                    int* restart_check)
{
  using namespace LAMMPS_NS;
  const int ntotal = nlocal + nghost;
  std::vector<double*> xp(ntotal), fp(ntotal);
  for (int i = 0; i < ntotal; i++) { xp[i] = x + 3 * (size_t)i; fp[i] = f + 3 * (size_t)i; }
  Atom atom = {xp.data(), fp.data(), type, nlocal, nghost, ntypes};
  Comm comm = {0};
  Error err;
  Neighbor neighbor;
  LAMMPS lmp = {&atom, &comm, &err, &neighbor};
  std::vector<int> ilist(nlocal);
  std::vector<int*> first(nlocal);
  std::vector<int> nn(numneigh, numneigh + nlocal);
  std::vector<int> flat(neigh_flat, neigh_flat + 0);
  size_t total = 0;
  for (int i = 0; i < nlocal; i++) total += numneigh[i];
  flat.assign(neigh_flat, neigh_flat + total);
  size_t off = 0;
  for (int i = 0; i < nlocal; i++) { ilist[i] = i; first[i] = flat.data() + off; off += numneigh[i]; }
  NeighList list = {nlocal, ilist.data(), nn.data(), first.data()};
  PairNN pair(&lmp);
  char buf[64];
  snprintf(buf, sizeof(buf), "%.17g", Rcut);
  char* sarg[1] = {buf};
  pair.settings(1, sarg);
  for (int i = 0; i < ntypes; i++) {
    char tbuf[16];
    snprintf(tbuf, sizeof(tbuf), "%d", coeff_types[i]);
    char* carg[2] = {tbuf, (char*)coeff_files[i]};
    pair.coeff(2, carg);
  }
  pair.init_style();
  if (!(neighbor.skin == 1.0 && neighbor.every == 10 && neighbor.delay == 0 && neighbor.requests[0]->full == 1 &&
        neighbor.requests[0]->half == 0 && neighbor.requests[0]->id == 1))
    return 2;
  double fdummy = 0;
  if (pair.single(0, 0, 1, 1, 1.0, 0.0, 0.0, fdummy) != 1) return 3;
  pair.init_list(1, &list);
  pair.eatom = eatom;
  pair.compute(3, 0);
  *eng_vdwl = pair.eng_vdwl;
  if (restart_check) {
    // write_restart -> a fresh PairNN -> read_restart -> compute: the energies must be the same numbers
    *restart_check = 1;
    FILE* rf = tmpfile();
    if (!rf) return 4;
    pair.write_restart(rf);
    rewind(rf);
    PairNN again(&lmp);
    again.read_restart(rf);
    fclose(rf);
    std::vector<double> f2(3 * (size_t)ntotal, 0.0), e2(nlocal, 0.0);
    std::vector<double*> fp2(ntotal);
    for (int i = 0; i < ntotal; i++) fp2[i] = f2.data() + 3 * (size_t)i;
    atom.f = fp2.data();
    again.init_list(1, &list);
    again.eatom = e2.data();
    again.compute(3, 0);
    atom.f = fp.data();
    if (again.eng_vdwl != pair.eng_vdwl) *restart_check = 0;
    for (int i = 0; i < nlocal; i++)
      if (e2[i] != eatom[i]) *restart_check = 0;
  }
  return 0;
}

}  // extern "C"
