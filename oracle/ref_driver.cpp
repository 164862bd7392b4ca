// Dump driver for the UNMODIFIED reference (biklooost/PROPhet), linked against the reference's own
// object files by oracle/build_ref.sh.  TEST INFRASTRUCTURE ONLY: nothing under prophet_b200/ may link
// or execute this.  It writes what the reference computes (neighbour lists, descriptors, energies,
// dE/dG, forces, gradients, optimizer steps) into a small tagged-binary file that
// tests/golden/make_goldens.py turns into the committed golden fixtures and that oracle/ checks its
// restatement against.
//
// This file is our own code; it only *calls* the reference through its classes.  It re-defines
// private/protected as public (after the standard headers) so it can read members the reference keeps private
// (Potential::nets/gradient/opt, Structure::trans_indices/translations, Neural_network::params).
//
//   ref_driver train <input_file> <out.bin> <n_iterations>
//       cwd holds input_file (+ PROPhet.xml / data dirs, + checkpoint files when checkpoint_in is set).
//       Follows main.cpp:50-63 (Setup -> System -> Potential) and restates the body of
//       Potential::train (Potential.cpp:607-672) so every per-iteration quantity can be dumped.
//   ref_driver md <md_case.txt> <out.bin>
//       Follows PairNN::settings/coeff/compute (pair_nn.cpp:499-526, 223-259, 112-188) without LAMMPS.
//   ref_driver optq <spec.txt> <out.bin>
//       spec.txt: P ncalls algorithm lambda_max ncycles threshold seed.  Drives the reference's Optimizer (through
//       dummy network nodes) on the quadratic f(x) = sum_i a_i (x_i - b_i)^2 and dumps x every 500 calls, the call at
//       which it converged and the next libc rand() value (the perturbation scheme draws random numbers).
//   ref_driver ks <samples.txt> <out.bin>
//       samples.txt: n1 v... n2 v...; calls Analysis::KS (Analysis.cpp:149-202) and dumps D and p.

#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <deque>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <limits>
#include <map>
#include <set>
#include <sstream>
#include <string>
#include <utility>
#include <vector>
#include <fenv.h>
#include <stddef.h>
#include <sys/stat.h>
#include <time.h>

// every standard header the reference's headers pull in is included above, so the two defines below
// only open up the reference's own classes
#define private public
#define protected public

#include "Setup.h"
#include "System.h"
#include "Potential.h"
#include "NN.h"
#include "Optimizer.h"
#include "Structure.h"
#include "Analysis.h"

using namespace std;

// ---- tagged binary writer:  "name dtype ndim d0 d1 ...\n" + raw little-endian payload -------------
static FILE* g_out = NULL;

static void put_raw(const string& name, char dtype, const vector<long>& dims, const void* data, size_t bytes)
{
  fprintf(g_out, "%s %c %d", name.c_str(), dtype, (int)dims.size());
  for (size_t i = 0; i < dims.size(); i++) fprintf(g_out, " %ld", dims[i]);
  fprintf(g_out, "\n");
  if (bytes) fwrite(data, 1, bytes, g_out);
}
static void put_d(const string& name, const vector<double>& v, vector<long> dims = vector<long>())
{
  if (dims.empty()) dims.push_back((long)v.size());
  put_raw(name, 'd', dims, v.empty() ? NULL : &v[0], v.size() * sizeof(double));
}
static void put_i(const string& name, const vector<int>& v, vector<long> dims = vector<long>())
{
  if (dims.empty()) dims.push_back((long)v.size());
  put_raw(name, 'i', dims, v.empty() ? NULL : &v[0], v.size() * sizeof(int));
}
static void put_scalar(const string& name, double v) { put_d(name, vector<double>(1, v)); }
static vector<long> dims2(long a, long b) { vector<long> d; d.push_back(a); d.push_back(b); return d; }
static string idx(const string& base, int i) { ostringstream s; s << base << "." << i; return s.str(); }

static void dump_rows(const string& name, const vector<vector<REAL> >& rows, int width)
{
  vector<double> flat;
  for (size_t i = 0; i < rows.size(); i++)
    for (int k = 0; k < width; k++) flat.push_back(k < (int)rows[i].size() ? rows[i][k] : 0.0);
  put_d(name, flat, dims2((long)rows.size(), width));
}

// ---------------------------------------------------------------------------------------------------
static int run_train(const string& input_file, int n_iter)
{
  Setup run_params(input_file);
  vector<System*> systems(run_params.Nsystems());
  for (int i = 0; i < run_params.Nsystems(); i++) systems.at(i) = new System(run_params.system(i), &run_params.F);

  // raw (pre-normalisation) coordinates as the readers produced them
  for (size_t s = 0; s < systems.size(); s++) {
    Structure& st = systems[s]->structure;
    vector<double> raw;
    for (size_t a = 0; a < st.pos.size(); a++) for (int k = 0; k < 3; k++) raw.push_back(st.pos[a][k]);
    put_d(idx("raw_pos", s), raw, dims2((long)st.pos.size(), 3));
    put_scalar(idx("raw_is_cartesian", s), st.CART ? 1.0 : 0.0);
  }

  Potential* pot = new Potential(systems, run_params.F);   // never deleted (mirrors the reference's own lifetime)

  put_i("atom_types", pot->atom_types);
  put_scalar("Rcut", pot->params.Rcut());
  put_scalar("n_systems", (double)systems.size());
  put_scalar("Ntrain", pot->Ntrain);
  put_scalar("Nval", pot->Nval);

  {
    Structure& st0 = systems[0]->structure;
    dump_rows("G1p", st0.G1p, 2);
    dump_rows("G2p", st0.G2p, 4);
    dump_rows("G3p", st0.G3p, 5);
    dump_rows("G4p", st0.G4p, 5);
    put_scalar("NG", st0.NG());
  }

  for (size_t s = 0; s < systems.size(); s++) {
    Structure& st = systems[s]->structure;
    int N = st.Natom;
    vector<double> cell;
    for (int k = 0; k < 3; k++) cell.push_back(st.a[k]);
    for (int k = 0; k < 3; k++) cell.push_back(st.b[k]);
    for (int k = 0; k < 3; k++) cell.push_back(st.c[k]);
    put_d(idx("cell", s), cell, dims2(3, 3));
    vector<double> pos; vector<int> Z;
    for (int a = 0; a < N; a++) {
      for (int k = 0; k < 3; k++) pos.push_back(st.pos[a][k]);
      Z.push_back(st.types[a].atomic_number());
    }
    put_d(idx("frac_pos", s), pos, dims2(N, 3));
    put_i(idx("Z", s), Z);
    put_scalar(idx("target", s), systems[s]->properties.target());
    int tag = systems[s]->train == "train" ? 0 : (systems[s]->train == "val" ? 1 : 2);
    put_scalar(idx("tag", s), tag);

    // neighbour list exactly as Structure::Set_Neighbor left it (Structure.cpp:1287-1381)
    vector<int> trans;
    for (size_t t = 0; t < st.translations.size(); t++) for (int k = 0; k < 3; k++) trans.push_back(st.translations[t][k]);
    put_i(idx("translations", s), trans, dims2((long)st.translations.size(), 3));
    put_i(idx("numneigh", s), st.numneigh);
    vector<int> nb_off(1, 0), nb_j, im_off(1, 0), im_t;
    for (int i = 0; i < N; i++) {
      for (size_t jj = 0; jj < st.firstneigh[i].size(); jj++) {
        nb_j.push_back(st.firstneigh[i][jj]);
        const vector<int>& T = st.trans_indices[i][jj];
        for (size_t tt = 0; tt < T.size(); tt++) im_t.push_back(T[tt]);
        im_off.push_back((int)im_t.size());
      }
      nb_off.push_back((int)nb_j.size());
    }
    put_i(idx("neigh_offset", s), nb_off);     // per atom -> range in neigh_j
    put_i(idx("neigh_j", s), nb_j);            // unique neighbour atom indices, ascending
    put_i(idx("image_offset", s), im_off);     // per (atom,neighbour) -> range in image_t
    put_i(idx("image_t", s), im_t);            // translation-table indices, ascending

    // descriptors G[NG][Natom] (Structure::init_G, Structure.cpp:874-892)
    int NG = st.NG();
    vector<double> G;
    for (int g = 0; g < NG; g++) for (int a = 0; a < N; a++) G.push_back(st.G[g][a]);
    put_d(idx("G", s), G, dims2(NG, N));
  }

  if (n_iter <= 0) return 0;

  // ---- Potential::train restated (Potential.cpp:607-672) ----
  pot->Precondition();
  for (map<int, Network*>::iterator it = pot->nets.begin(); it != pot->nets.end(); ++it) {
    Neural_network* nn = dynamic_cast<Neural_network*>(it->second);
    put_d(idx("input_mean", it->first), nn->params.input_mean);
    put_d(idx("input_scale", it->first), nn->params.input_variance);
    put_i(idx("layers", it->first), nn->NNodes);
  }
  put_scalar("output_mean", pot->output_mean);
  put_scalar("output_variance", pot->output_variance);

  pot->opt = new Optimizer(pot->mpi, pot->params, pot->Ntrain);
  pot->opt->set_early_stop(pot->early_stop);
  pot->init_optimizer();
  pot->syncronize();

  // per-atom energies with the initial parameters (Neural_network::evaluate_MD(int), NN.cpp:585-612)
  for (size_t s = 0; s < systems.size(); s++) {
    vector<double> e_atom;
    for (int a = 0; a < systems[s]->structure.Natom; a++)
      e_atom.push_back(pot->nets[systems[s]->structure.types[a].atomic_number()]->evaluate_MD((int)s));
    put_d(idx("E_atom", s), e_atom);
  }

  for (int iter = 0; iter < n_iter && !pot->opt->is_converged(); iter++) {
    put_d(idx("x_before", iter), pot->opt->get_parameters());
    REAL SSE = 0, vSSE = 0;
    int nval = 0;
    pot->gradient.assign(pot->Ntotal_params, 0.0);
    vector<double> E_struct(systems.size(), 0.0);
    for (size_t s = 0; s < systems.size(); s++) {
      if (systems[s]->train == "train") {
        REAL output = 0;
        for (int a = 0; a < systems[s]->structure.Natom; a++)
          output += pot->nets[systems[s]->structure.types[a].atomic_number()]->train_MD((int)s);
        REAL Error = (output + pot->output_mean) - systems[s]->properties.target();
        E_struct[s] = output + pot->output_mean;
        SSE += Error * Error;
        for (map<int, Network*>::iterator it = pot->nets.begin(); it != pot->nets.end(); ++it) {
          const vector<REAL> g = it->second->get_gradient();
          vector<REAL>::iterator v = pot->indices[it->first];
          for (size_t i = 0; i < g.size(); i++) { (*v) += 2 * Error * g[i]; ++v; }
        }
      } else if (systems[s]->train == "val") {
        REAL output = 0;
        for (int a = 0; a < systems[s]->structure.Natom; a++)
          output += pot->nets[systems[s]->structure.types[a].atomic_number()]->train_MD((int)s);
        REAL Error = (output + pot->output_mean) - systems[s]->properties.target();
        E_struct[s] = output + pot->output_mean;
        vSSE += Error * Error;
        nval++;
      }
    }
    put_d(idx("E_struct", iter), E_struct);
    put_scalar(idx("SSE", iter), SSE);
    put_scalar(idx("vSSE", iter), vSSE);
    put_scalar(idx("nval", iter), nval);
    put_d(idx("gradient", iter), pot->gradient);
    pot->opt->set_val_sse(vSSE, nval);
    pot->opt->update_network(SSE, pot->gradient);
    put_d(idx("x_after", iter), pot->opt->get_parameters());
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// md case file (text):
//   Rcut <r>
//   ntypes <n>
//   coeff <lammps_type> <checkpoint_file>      (n lines, in pair_coeff order)
//   atoms <nlocal> <nghost>
//   <type> <x> <y> <z>                         (nlocal+nghost lines)
//   neigh <i> <count> j0 j1 ...                (nlocal lines, full list incl. ghosts)
static int run_md(const string& case_file)
{
  ifstream in(case_file.c_str());
  if (!in.is_open()) { fprintf(stderr, "cannot open %s\n", case_file.c_str()); return 2; }
  string key;
  double Rcut = 0; int ntypes = 0, nlocal = 0, nghost = 0;
  vector<pair<int, string> > coeffs;
  in >> key >> Rcut;
  in >> key >> ntypes;
  for (int i = 0; i < ntypes; i++) { int t; string f; in >> key >> t >> f; coeffs.push_back(make_pair(t, f)); }
  in >> key >> nlocal >> nghost;
  int ntotal = nlocal + nghost;
  vector<double> xs(3 * ntotal), fs(3 * ntotal, 0.0);
  vector<int> type(ntotal);
  vector<double*> x(ntotal), f(ntotal);
  for (int i = 0; i < ntotal; i++) {
    in >> type[i] >> xs[3 * i] >> xs[3 * i + 1] >> xs[3 * i + 2];
    x[i] = &xs[3 * i]; f[i] = &fs[3 * i];
  }
  vector<vector<int> > neigh(nlocal);
  for (int i = 0; i < nlocal; i++) {
    int ii, cnt; in >> key >> ii >> cnt;
    neigh[ii].resize(cnt);
    for (int k = 0; k < cnt; k++) in >> neigh[ii][k];
  }
  if (!in) { fprintf(stderr, "md case file truncated\n"); return 2; }

  // PairNN::settings (pair_nn.cpp:499-526)
  Functional_params params;
  params.Rcut(Rcut);
  params.is_MD = true;
  vector<string> inputs(1, "structure");
  params.set_inputs(inputs);
  params.set_output("energy");
  System* sysdata = new System();
  sysdata->structure.set_structure(&x[0], &type[0], nlocal, nghost);
  sysdata->properties.lock(true);
  sysdata->structure.periodic = 0;
  Potential* potential = new Potential(params);    // never deleted: its mpi pointer is uninitialised
  potential->add_system(sysdata);
  // PairNN::coeff (pair_nn.cpp:245-246)
  for (size_t c = 0; c < coeffs.size(); c++) {
    potential->insert_atom_type(coeffs[c].first, (char*)coeffs[c].second.c_str(), sysdata);
    sysdata->structure.lammps_conv[coeffs[c].first] = potential->ret_params().current_atom_type;
  }
  // PairNN::compute (pair_nn.cpp:144-182)
  Structure& st = sysdata->structure;
  st.inum = nlocal;
  st.ilist.resize(nlocal);
  st.numneigh.resize(nlocal);
  for (int i = 0; i < nlocal; i++) { st.ilist[i] = i; st.numneigh[i] = (int)neigh[i].size(); }
  st.firstneigh.assign(nlocal, vector<int>(1, 0));
  st.set_structure(&x[0], &type[0], nlocal, nghost);
  for (int i = 0; i < nlocal; i++) st.firstneigh[i] = neigh[i];
  st.periodic = 0;

  vector<vector<REAL> > dE_dG(nlocal, vector<REAL>(1, 0.0));
  vector<double> eatom(nlocal, 0.0), Gdump;
  double etotal = 0;
  int NG = 0;
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (int ii = 0; ii < nlocal; ii++) {
    double E = potential->evaluate_MD(ii, type[ii], dE_dG[ii]);
    eatom[ii] = E;
    etotal += E;
    NG = st.NG();
    for (int g = 0; g < NG; g++) Gdump.push_back(st.G[g][0]);
  }
  etotal += potential->energy_shift();
  st.Get_Forces(dE_dG, &f[0]);
  clock_gettime(CLOCK_MONOTONIC, &t1);
  // wall time of the per-step work of PairNN::compute (descriptors + networks + forces), for bench.py's CPU arm
  printf("MD_COMPUTE_SECONDS %.6f\n", (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec));

  put_scalar("NG", NG);
  put_d("G", Gdump, dims2(nlocal, NG));
  put_d("eatom", eatom);
  put_scalar("etotal", etotal);
  put_scalar("energy_shift", potential->energy_shift());
  vector<double> dedg;
  for (int i = 0; i < nlocal; i++) for (int g = 0; g < NG; g++) dedg.push_back(dE_dG[i][g]);
  put_d("dE_dG", dedg, dims2(nlocal, NG));
  put_d("f", fs, dims2(ntotal, 3));
  return 0;
}

static int run_optq(const string& file)
{
  ifstream in(file.c_str());
  if (!in.is_open()) { fprintf(stderr, "cannot open %s\n", file.c_str()); return 2; }
  int P = 0, ncalls = 0, ncycles = 1, seed = 1;
  string algorithm;
  double lambda_max = 0, threshold = 0;
  in >> P >> ncalls >> algorithm >> lambda_max >> ncycles >> threshold >> seed;
  srand(seed);
  Functional_params F;
  F.my_training_algorithm = algorithm;
  F.my_lambda_max = lambda_max;
  F.my_Ncycles = ncycles;
  F.my_threshold = threshold;
  F.my_Niterations = 1 << 30;
  F.my_Nprint = 1 << 30;
  Parallel* mpi = new Parallel();
  // one dummy node with P-1 inputs holds the P parameters
  vector<Network_node*> nodes(1, new Neural_network_node(P - 1, "linear"));
  Optimizer* opt = new Optimizer(mpi, F, 1);
  opt->init(nodes);
  opt->internal_init();
  vector<REAL> a(P), b(P);
  for (int i = 0; i < P; i++) { a[i] = 0.5 + 0.37 * ((i * 7) % 5); b[i] = 0.25 * ((i * 3) % 7) - 0.6; }
  put_d("x0", opt->get_parameters());
  int converged_at = -1;
  for (int call = 0; call < ncalls; call++) {
    vector<REAL> x = opt->get_parameters();
    REAL f = 0;
    vector<REAL> g(P);
    for (int i = 0; i < P; i++) { const REAL d = x[i] - b[i]; f += a[i] * d * d; g[i] = 2 * a[i] * d; }
    opt->update_network(f, g);
    if ((call + 1) % 500 == 0) put_d(idx("x", call + 1), opt->get_parameters());
    if (opt->is_converged()) { converged_at = call + 1; break; }
  }
  put_d("x_final", opt->get_parameters());
  put_scalar("converged_at", converged_at);
  put_scalar("next_rand", (double)rand());
  return 0;
}

static int run_ks(const string& file)
{
  ifstream in(file.c_str());
  if (!in.is_open()) { fprintf(stderr, "cannot open %s\n", file.c_str()); return 2; }
  int n1 = 0, n2 = 0;
  in >> n1;
  vector<REAL> a(n1);
  for (int i = 0; i < n1; i++) in >> a[i];
  in >> n2;
  vector<REAL> b(n2);
  for (int i = 0; i < n2; i++) in >> b[i];
  Analysis A;
  REAL D = A.KS(a, b);
  put_scalar("D", D);
  put_scalar("p", A.p);
  return 0;
}

int main(int argc, char** argv)
{
  if (argc < 4) {
    fprintf(stderr, "usage: ref_driver train <input_file> <out.bin> <n_iter> | md <case.txt> <out.bin> | ks <samples.txt> <out.bin>\n");
    return 2;
  }
  string mode = argv[1];
  g_out = fopen(argv[3], "wb");
  if (!g_out) { fprintf(stderr, "cannot write %s\n", argv[3]); return 2; }
  int rc = 2;
  if (mode == "train") rc = run_train(argv[2], argc > 4 ? atoi(argv[4]) : 0);
  else if (mode == "md") rc = run_md(argv[2]);
  else if (mode == "ks") rc = run_ks(argv[2]);
  else if (mode == "optq") rc = run_optq(argv[2]);
  fclose(g_out);
  return rc;
}
