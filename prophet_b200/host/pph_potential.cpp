// HostNet, Parallel and Potential of the host layer (see pph.h for the reference map).  All arithmetic of the
// hot path happens behind the C ABI on the GPU; this file moves parameters, files and tables.
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <set>
#include <sstream>

#include <sys/stat.h>
#include <ctime>

#include "pph.h"

namespace pph {

// ---- transfer functions (names: NN_node.cpp:120-171) -----------------------------------------------------------
int transfer_id(const std::string& name)
{
  static const char* names[] = {"linear", "tanh", "tanh_spline", "tanh_crazy", "cubic", "pulse", "sin", "periodic_tanh", "relu"};
  for (int i = 0; i < 9; i++)
    if (name == names[i]) return i;
  PPH_ERROR("Unrecognized transfer function '" + name + "'");
}

// ---- HostNet --------------------------------------------------------------------------------------------------
static REAL rand_uniform() { return (REAL)rand() / RAND_MAX; }
static REAL rand_normal(REAL mean, REAL std_dev)
{
  // Common.h:68-75 (the reference leaves the order of its two rand() calls to the compiler; fixed here)
  REAL u1 = rand_uniform();
  REAL u2 = rand_uniform();
  return mean + std_dev * std::sqrt(-2.0 * std::log(u1)) * std::cos(2.0 * 3.141592653589793 * u2);
}

void HostNet::init_random()
{
  // Neural_network ctor + Neural_network_node ctor (NN.cpp:51-80, NN_node.cpp:52-62)
  weights.clear();
  node_transfer.clear();
  const int L = (int)sizes.size() - 1;
  for (int l = 0; l < L; l++) {
    for (int n = 0; n < sizes[l + 1]; n++) {
      for (int i = 0; i < sizes[l]; i++) weights.push_back(rand_normal(0, 10.5 / std::sqrt((REAL)sizes[l])));
      weights.push_back(rand_normal(0, 1));
      node_transfer.push_back(l == L - 1 ? "linear" : params.transfer_function);
    }
  }
}

void HostNet::load(std::istream& in)
{
  // Neural_network::load (NN.cpp:158-200)
  weights.clear();
  node_transfer.clear();
  std::string line, junk, tf;
  const int L = (int)sizes.size() - 1;
  for (int l = 0; l < L; l++) {
    std::getline(in, line);
    while (!in.eof() && line.find("[[ layer ") == std::string::npos) std::getline(in, line);
    for (int n = 0; n < sizes[l + 1]; n++) {
      std::getline(in, line);
      while (!in.eof() && line.find("[ node ") == std::string::npos) std::getline(in, line);
      std::istringstream Line(line);
      Line >> junk >> junk >> junk >> junk >> tf;
      node_transfer.push_back(tf);
      transfer_id(tf);
      std::getline(in, line);
      std::istringstream W(line);
      REAL v;
      int cnt = 0;
      while (W >> v) { weights.push_back(v); cnt++; }
      if (cnt != sizes[l]) PPH_ERROR("Parameters not set correctly");
      std::getline(in, line);
      std::istringstream B(line);
      B >> v;
      weights.push_back(v);
    }
  }
}

void HostNet::print(std::ostream& out) const
{
  // Neural_network::print (NN.cpp:828-847) + Neural_network_node::print (NN_node.cpp:232-243)
  params.print(out);
  const int L = (int)sizes.size() - 1;
  int node = 0;
  size_t p = 0;
  for (int l = 0; l < L; l++) {
    out << "[[ layer " << l << " ]] " << std::endl;
    const REAL scale = (l == 0 || l == L - 1) ? 1.0 : params.dropout;
    for (int n = 0; n < sizes[l + 1]; n++) {
      out << "  [ node " << node << " ]  " << node_transfer[node] << std::endl;
      for (int i = 0; i < sizes[l]; i++) out << "   " << std::setprecision(18) << scale * weights[p + i] << "  ";
      out << std::endl;
      out << "   " << std::setprecision(18) << scale * weights[p + sizes[l]] << std::endl;
      p += sizes[l] + 1;
      node++;
    }
  }
}

// ---- Parallel -------------------------------------------------------------------------------------------------
static int env_int(const char* a, const char* b, int dflt)
{
  const char* v = getenv(a);
  if (!v && b) v = getenv(b);
  return v ? atoi(v) : dflt;
}

Parallel::Parallel() : ctx(NULL)
{
  my_rank = env_int("PP_RANK", "RANK", 0);
  my_nprocs = env_int("PP_WORLD_SIZE", "WORLD_SIZE", 1);
  my_local_rank = env_int("PP_LOCAL_RANK", "LOCAL_RANK", my_rank);
  if (my_nprocs < 1 || my_rank < 0 || my_rank >= my_nprocs) PPH_ERROR("bad RANK / WORLD_SIZE in the environment");
}

void Parallel::attach(pp_context* c)
{
  ctx = c;
  if (my_nprocs == 1) return;
  // rank 0 publishes the ncclUniqueId in a rendezvous file; the others wait for it
  std::string path = getenv("PP_RENDEZVOUS") ? getenv("PP_RENDEZVOUS") : "";
  if (path.empty()) {
    std::ostringstream s;
    s << "/tmp/prophet_b200_" << (getenv("MASTER_PORT") ? getenv("MASTER_PORT") : "0") << ".id";
    path = s.str();
  }
  unsigned char id[128];
  if (my_rank == 0) {
    unlink(path.c_str());      // a file left behind by a run that died must not be taken for this run's
    PPH_CALL(pp_comm_unique_id(id));
    std::string tmp = path + ".tmp";
    std::ofstream f(tmp.c_str(), std::ios::binary);
    f.write((const char*)id, 128);
    f.close();
    rename(tmp.c_str(), path.c_str());
  } else {
    for (int tries = 0;; tries++) {
      // only a file written recently counts: one older than ten minutes is the leftover of an earlier run that rank 0
      // has not yet removed
      struct stat sb;
      const bool fresh = stat(path.c_str(), &sb) == 0 && difftime(time(NULL), sb.st_mtime) < 600.0;
      std::ifstream f(path.c_str(), std::ios::binary);
      if (fresh && f.is_open() && f.read((char*)id, 128)) break;
      if (tries > 6000) PPH_ERROR("timed out waiting for the rendezvous file " + path);
      usleep(10000);
    }
  }
  PPH_CALL(pp_comm_init(ctx, my_nprocs, my_rank, id));
  // everyone has joined once a first collective completes; rank 0 may then remove the file
  std::vector<REAL> one(1, 1.0);
  allreduce(one);
  if (my_rank == 0) unlink(path.c_str());
}

void Parallel::allreduce(std::vector<REAL>& v)
{
  if (my_nprocs == 1 || !ctx || v.empty()) return;
  PPH_CALL(pp_comm_allreduce_host(ctx, v.data(), (int)v.size()));
}

REAL Parallel::allreduce(REAL v)
{
  std::vector<REAL> t(1, v);
  allreduce(t);
  return t[0];
}

// ---- Potential ------------------------------------------------------------------------------------------------
static int tag_id(const std::string& t) { return t == "train" ? PP_TAG_TRAIN : (t == "val" ? PP_TAG_VAL : PP_TAG_OTHER); }

Potential::Potential(const Setup& setup, Parallel* mpi_in) : mpi(mpi_in), params(setup.F), ctx(NULL)
{
  targets_shifted = (setup.F.output == "energy" && !setup.F.FE.empty());
  // Potential::Potential(systems, F) (Potential.cpp:60-205)
  Nlocal = setup.Nsystems();
  PPH_CALL(pp_create(mpi->local_rank(), 64, &ctx));
  mpi->attach(ctx);

  std::vector<StructureData> structs(Nlocal);
  std::set<int> zs;
  for (int i = 0; i < Nlocal; i++) {
    structs[i] = setup.load(i);
    for (size_t a = 0; a < structs[i].Z.size(); a++) zs.insert(structs[i].Z[a]);
  }
  // union of atom types over ranks (get_all_atom_types, Potential.cpp:888-917): presence flags summed over ranks
  {
    std::vector<REAL> present(120, 0.0);
    for (std::set<int>::iterator it = zs.begin(); it != zs.end(); ++it) present[*it] = 1.0;
    mpi->allreduce(present);
    atom_types.clear();
    for (int z = 1; z < 120; z++)
      if (present[z] > 0) atom_types.push_back(z);
  }
  if (atom_types.empty()) PPH_ERROR("no atoms found in the given systems");
  params.atomic_numbers = atom_types;
  params.atomic_symbols.clear();
  for (size_t i = 0; i < atom_types.size(); i++) params.atomic_symbols.push_back(atomic_symbol(atom_types[i]));
  Nsystems = (int)std::lround(mpi->allreduce((REAL)Nlocal));

  if (!params.input_file.empty() && mpi->io_node()) std::cout << "Reading from checkpoint files:" << std::endl << std::endl;
  Ntotal_params = 0;
  for (size_t e = 0; e < atom_types.size(); e++) {
    const int Z = atom_types[e];
    HostNet net;
    net.Z = Z;
    net.preconditioned = false;
    std::string filename;
    if (!params.input_file.empty()) {
      filename = params.input_file + "_" + atomic_symbol(Z);
      std::ifstream file(filename.c_str());
      if (!file.is_open()) PPH_ERROR("Found atom type " + atomic_symbol(Z) + " but could not open restart file '" + filename + "'");
      if (mpi->io_node()) std::cout << "Reading parameters for " + atomic_symbol(Z) + " from file '" + filename + "'" << std::endl;
      params.read(file);      // rows, means, hidden ... now come from the file (Potential.cpp:96)
      if (mpi->io_node()) {
        std::cout << "Hidden layers :  ";
        for (size_t l = 0; l < params.hidden.size(); l++) std::cout << params.hidden[l] << "  ";
        std::cout << std::endl;
        if (!params.FE.empty()) {                                             // Potential.cpp:103-108
          std::cout.precision(9);
          std::cout << params.current_atom_type << " Free Energy : " << params.FE[params.current_atom_type] << std::endl;
          std::cout.precision(6);
        }
        std::cout << std::endl;
      }
      // the reference keeps the union found in the data, not the file's own list
      params.atomic_numbers = atom_types;
      params.atomic_symbols.clear();
      for (size_t i = 0; i < atom_types.size(); i++) params.atomic_symbols.push_back(atomic_symbol(atom_types[i]));
      net.preconditioned = !params.input_mean.empty();
    }
    if (e == 0) {
      // first element: descriptors of every local system (Potential.cpp:113-150, Structure::init_G)
      params.auto_rows();
      upload_descriptor_rows();
      std::map<int, int> zidx;
      for (size_t i = 0; i < atom_types.size(); i++) zidx[atom_types[i]] = (int)i;
      Ntrain = Nval = Nother = 0;
      const int BATCH = 512;
      for (int b0 = 0; b0 < Nlocal; b0 += BATCH) {
        const int b1 = std::min(Nlocal, b0 + BATCH);
        std::vector<int> nat, elem, tg;
        std::vector<double> cells, frac, tgt;
        for (int i = b0; i < b1; i++) {
          const StructureData& s = structs[i];
          nat.push_back(s.natoms());
          cells.insert(cells.end(), s.cell, s.cell + 9);
          frac.insert(frac.end(), s.pos.begin(), s.pos.end());
          for (size_t a = 0; a < s.Z.size(); a++) elem.push_back(zidx[s.Z[a]]);
          tgt.push_back(s.target);
          tg.push_back(tag_id(s.train));
          targets.push_back(s.target);
          tags.push_back(s.train);
          natoms.push_back(s.natoms());
          atom_Z.push_back(std::vector<unsigned char>(s.Z.begin(), s.Z.end()));
          if (s.train == "train") Ntrain++;
          else if (s.train == "val") Nval++;
          else Nother++;
        }
        PPH_CALL(pp_add_structures(ctx, b1 - b0, nat.data(), cells.data(), frac.data(), elem.data(), tgt.data(), tg.data()));
      }
      if (params.print_gvectors) {
        // print_gvector = true (Potential.cpp:114-136): "<rank>.gvector", one line of NG comma-terminated values per
        // atom, a blank line after every structure, stream default precision
        std::ostringstream name;
        name << mpi->rank() << ".gvector";
        std::ofstream gv(name.str().c_str());
        const int NG = pp_num_descriptors(ctx);
        std::vector<double> G;
        for (int s = 0; s < Nlocal; s++) {
          const int n = structs[s].natoms();
          G.assign((size_t)NG * n, 0.0);
          PPH_CALL(pp_get_descriptors(ctx, s, G.data()));
          for (int a = 0; a < n; a++) {
            for (int k = 0; k < NG; k++) gv << G[(size_t)k * n + a] << ",";
            gv << "\n";
          }
          gv << "\n";
        }
      }
    }
    params.current_atom_type = atomic_symbol(Z);
    net.params = params;
    net.sizes.clear();
    net.sizes.push_back(pp_num_descriptors(ctx));
    for (size_t l = 0; l < params.hidden.size(); l++) net.sizes.push_back(params.hidden[l]);
    net.sizes.push_back(1);
    // the reference's node constructors draw random weights even when a checkpoint then overwrites them
    // (NN.cpp:51-80, NN_node.cpp:52-62): drawing here as well keeps the libc rand() stream aligned for what
    // consumes it later (-optimize, the SGD shuffle)
    net.init_random();
    if (!params.input_file.empty()) {
      std::ifstream file(filename.c_str());
      net.load(file);
    }
    if ((int)net.params.input_mean.size() != net.sizes[0]) {
      net.params.input_mean.assign(net.sizes[0], 0.0);
      net.params.input_variance.assign(net.sizes[0], 1.0);
    }
    Ntotal_params += net.Nparameters();
    nets[Z] = net;
  }
  Ntrain = (int)std::lround(mpi->allreduce((REAL)Ntrain));
  Nval = (int)std::lround(mpi->allreduce((REAL)Nval));
  Nother = (int)std::lround(mpi->allreduce((REAL)Nother));
  early_stop = Nval > 0;
  if (mpi->io_node()) {
    std::cout << "Data Info: \n";
    std::cout << "Train: " << Ntrain << " Val: " << Nval << " Other: " << Nother << std::endl << std::endl;
    double ratio = (double)Ntrain / Ntotal_params;
    if (ratio <= 1) std::cout << "<< WARNING >> ";
    std::cout << "Ratio of training data to parameters =  " << ratio << std::endl << std::endl;
  }
  output_mean = params.output_mean;
  output_variance = params.output_variance;
  upload_networks();
}

Potential::~Potential()
{
  if (ctx) pp_destroy(ctx);
}

void Potential::upload_networks()
{
  for (size_t e = 0; e < atom_types.size(); e++) {
    HostNet& n = nets[atom_types[e]];
    std::vector<int> tf;
    for (size_t k = 0; k < n.node_transfer.size(); k++) tf.push_back(transfer_id(n.node_transfer[k]));
    PPH_CALL(pp_set_network(ctx, (int)e, (int)n.sizes.size() - 1, n.sizes.data(), tf.data(), n.weights.data(),
                            n.params.input_mean.data(), n.params.input_variance.data()));
  }
  PPH_CALL(pp_set_output_scaling(ctx, output_mean, output_variance));
}

std::vector<REAL> Potential::all_parameters() const
{
  std::vector<REAL> x;
  for (size_t e = 0; e < atom_types.size(); e++) {
    const HostNet& n = nets.find(atom_types[e])->second;
    x.insert(x.end(), n.weights.begin(), n.weights.end());
  }
  return x;
}

void Potential::set_all_parameters(const std::vector<REAL>& x)
{
  size_t p = 0;
  for (size_t e = 0; e < atom_types.size(); e++) {
    HostNet& n = nets[atom_types[e]];
    std::copy(x.begin() + p, x.begin() + p + n.weights.size(), n.weights.begin());
    p += n.weights.size();
  }
}

void Potential::upload_descriptor_rows()
{
  std::vector<double> g1, g2, g3, g4;
  for (size_t r = 0; r < params.G1.size(); r++) g1.insert(g1.end(), params.G1[r].begin(), params.G1[r].end());
  for (size_t r = 0; r < params.G2.size(); r++) g2.insert(g2.end(), params.G2[r].begin(), params.G2[r].end());
  for (size_t r = 0; r < params.G3.size(); r++) g3.insert(g3.end(), params.G3[r].begin(), params.G3[r].end());
  for (size_t r = 0; r < params.G4.size(); r++) g4.insert(g4.end(), params.G4[r].begin(), params.G4[r].end());
  PPH_CALL(pp_set_descriptors(ctx, (int)atom_types.size(), (int)params.G1.size(), g1.data(), (int)params.G2.size(), g2.data(),
                              (int)params.G3.size(), g3.data(), (int)params.G4.size(), g4.data(), params.Rcut));
}

void Potential::Precondition()
{
  // Potential.cpp:387-502
  if (mpi->io_node()) std::cout << "Preconditioning ... " << std::endl << std::endl;
  bool any = false;
  for (size_t e = 0; e < atom_types.size(); e++)
    if (!nets[atom_types[e]].preconditioned) any = true;
  if (any) {
    PPH_CALL(pp_precondition(ctx, params.input_precondition ? 1 : 0));
    for (size_t e = 0; e < atom_types.size(); e++) {
      HostNet& n = nets[atom_types[e]];
      if (n.preconditioned) {
        // means loaded from a checkpoint win (NN.h:100-103): put them back
        PPH_CALL(pp_set_input_scaling(ctx, (int)e, n.params.input_mean.data(), n.params.input_variance.data()));
      } else {
        PPH_CALL(pp_get_input_scaling(ctx, (int)e, n.params.input_mean.data(), n.params.input_variance.data()));
        n.preconditioned = true;      // set_means makes is_preconditioned() true: a later train() keeps these (NN.h:92-103)
      }
    }
  }
  if (params.output_precondition) {
    // mean / variance of the training targets over all ranks (Potential.cpp:461-495)
    std::vector<REAL> acc(2, 0.0);
    for (int i = 0; i < Nlocal; i++)
      if (tags[i] == "train") { acc[0] += targets[i]; acc[1] += 1.0; }
    mpi->allreduce(acc);
    output_mean = acc[1] > 0 ? acc[0] / acc[1] : 0.0;
    REAL var = 0.0;
    for (int i = 0; i < Nlocal; i++)
      if (tags[i] == "train") var += (targets[i] - output_mean) * (targets[i] - output_mean);
    var = mpi->allreduce(var);
    output_variance = acc[1] > 0 ? var / acc[1] : 1.0;
  } else {
    output_mean = 0;
    output_variance = 1;
  }
  for (size_t e = 0; e < atom_types.size(); e++) {
    nets[atom_types[e]].params.output_mean = output_mean;
    nets[atom_types[e]].params.output_variance = output_variance;
  }
  PPH_CALL(pp_set_output_scaling(ctx, output_mean, output_variance));
}

REAL Potential::train()
{
  // Potential.cpp:607-693
  Precondition();
  Optimizer opt(mpi, params, Ntrain);
  opt.set_early_stop(early_stop);
  opt.init(all_parameters());
  std::vector<REAL> gradient(Ntotal_params, 0.0);
  while (!opt.is_converged()) {
    double SSE = 0, vSSE = 0;
    long long nval = 0;
    // the loop over systems and atoms, the per-structure error, the gradient assembly and the Reduce of
    // Optimizer::update_network all happen inside this one call
    // `sgd = true` only shuffles the order in which the reference visits its systems within a full-batch pass
    // (Potential.cpp:632-634): the sums are the same up to rounding, but std::random_shuffle draws Nlocal - 1 numbers
    // from libc rand() per epoch, which whatever reads that stream later (-optimize) must see
    if (params.SGD)
      for (int k = 1; k < Nlocal; k++) (void)rand();
    PPH_CALL(pp_epoch(ctx, opt.parameters().data(), gradient.data(), &SSE, &vSSE, &nval));
    opt.set_val_sse(vSSE, nval);
    opt.update_network(SSE, gradient);
    set_all_parameters(opt.parameters());
    if (opt.is_checkpoint()) save();
    if (opt.is_backup()) backup(opt.iteration());
  }
  save();
  return std::sqrt(opt.get_error());
}

// Redraw until the value falls in (lo, hi] (open_lo) or [lo, hi]
static REAL perturbed(REAL centre, REAL sd, REAL lo, bool open_lo, REAL hi)
{
  for (;;) {
    REAL v = rand_normal(0, sd) + centre;
    if ((open_lo ? v > lo : v >= lo) && v <= hi) return v;
  }
}

void Potential::optimize_Gs()
{
  // Potential.cpp:709-877 (J. Chem. Phys. 144, 224103): simulated annealing over eta / Rs (G2), eta / xi (G3, G4).
  // Every step perturbs the LAST ACCEPTED rows, recomputes all descriptors on the GPU (one launch over the resident
  // structures), retrains from the current weights and accepts with the Metropolis rule on the final RMS error.
  // Random numbers are drawn on rank 0 only and the rows broadcast, as the reference does.
  const REAL max_eta = 65;
  std::vector<std::vector<REAL> > G1 = params.G1, G2 = params.G2, G3 = params.G3, G4 = params.G4;
  REAL T = params.T0, last_E = 9e9;
  int Naccepted = 0, counter = 0;
  std::ofstream log_file;
  log_file.open("logfile.out");
  if (mpi->io_node()) {
    log_file << "  T    accepted    Error" << std::endl;
    log_file << " -------------------------" << std::endl;
  }
  while (T >= params.Tf) {
    if (mpi->io_node()) {
      Naccepted = 0;
      log_file << T << " : ";
    }
    for (int Step = 1; Step <= params.Nsteps_per_T; Step++) {
      if (mpi->io_node()) std::cout << "Step : " << Step << std::endl;
      if (counter) {
        if (mpi->io_node()) {
          for (size_t i = 0; i < params.G2.size(); i++) {
            params.G2[i][2] = perturbed(G2[i][2], params.step_size, 0, true, max_eta);
            params.G2[i][3] = perturbed(G2[i][3], params.step_size, 0, false, params.Rcut);
          }
          for (size_t i = 0; i < params.G3.size(); i++) {
            params.G3[i][2] = perturbed(G3[i][2], params.step_size, 0, true, max_eta);
            params.G3[i][3] = perturbed(G3[i][3], params.step_size, 0, false, params.Rcut);
          }
          for (size_t i = 0; i < params.G4.size(); i++) {
            params.G4[i][2] = perturbed(G4[i][2], params.step_size, 0, true, max_eta);
            params.G4[i][3] = perturbed(G4[i][3], params.step_size, 0, true, max_eta);
          }
        }
        // Bcast of the rows: zero everywhere but rank 0, then sum
        std::vector<REAL> flat;
        for (size_t i = 0; i < params.G2.size(); i++) { flat.push_back(params.G2[i][2]); flat.push_back(params.G2[i][3]); }
        for (size_t i = 0; i < params.G3.size(); i++) { flat.push_back(params.G3[i][2]); flat.push_back(params.G3[i][3]); }
        for (size_t i = 0; i < params.G4.size(); i++) { flat.push_back(params.G4[i][2]); flat.push_back(params.G4[i][3]); }
        if (mpi->Nprocs() > 1) {
          if (!mpi->io_node()) std::fill(flat.begin(), flat.end(), 0.0);
          mpi->allreduce(flat);
        }
        size_t k = 0;
        for (size_t i = 0; i < params.G2.size(); i++) { params.G2[i][2] = flat[k++]; params.G2[i][3] = flat[k++]; }
        for (size_t i = 0; i < params.G3.size(); i++) { params.G3[i][2] = flat[k++]; params.G3[i][3] = flat[k++]; }
        for (size_t i = 0; i < params.G4.size(); i++) { params.G4[i][2] = flat[k++]; params.G4[i][3] = flat[k++]; }
        upload_descriptor_rows();
        PPH_CALL(pp_recompute_descriptors(ctx));
      }
      REAL Error = train();
      if (mpi->io_node()) {
        std::cout << "Error =  " << Error << std::endl;
        if (rand_uniform() < std::exp(-(Error - last_E) / T)) {
          G2 = params.G2;
          G3 = params.G3;
          G4 = params.G4;
          last_E = Error;
          Naccepted++;
          std::ofstream pf("G_parameters.out");
          const std::vector<std::vector<REAL> >* sets[4] = {&G1, &G2, &G3, &G4};
          const char* names[4] = {"G1  ", "G2  ", "G3  ", "G4  "};
          for (int g = 0; g < 4; g++)
            for (size_t i = 0; i < sets[g]->size(); i++) {
              if (g == 0) pf << std::setprecision(10);      // only a G1 row raises the precision (Potential.cpp:826-827)
              pf << names[g];
              for (size_t j = 0; j < (*sets[g])[i].size(); j++) pf << (*sets[g])[i][j] << " ";
              pf << std::endl;
            }
        }
      }
      counter++;
    }
    if (mpi->io_node()) log_file << (double)Naccepted / (double)params.Nsteps_per_T << "  " << last_E << std::endl;
    T *= params.dT;
  }
}

void Potential::save()
{
  if (!mpi->io_node()) return;
  for (size_t e = 0; e < atom_types.size(); e++) {
    const HostNet& n = nets[atom_types[e]];
    std::string filename = params.output_file + "_" + atomic_symbol(n.Z);
    std::ofstream out(filename.c_str());
    n.print(out);
  }
}

void Potential::backup(int niter)
{
  if (!mpi->io_node()) return;
  for (size_t e = 0; e < atom_types.size(); e++) {
    const HostNet& n = nets[atom_types[e]];
    std::ostringstream name;
    name << params.output_file << "_" << niter - 1 << "_" << atomic_symbol(n.Z);
    std::ofstream out(name.str().c_str());
    n.print(out);
  }
}

// gather per-system values of all ranks into the original input order (create_system_map, Potential.cpp:1179-1203):
// system g lives on rank g % Nprocs at local position g / Nprocs
static std::vector<REAL> gather_in_order(Parallel* mpi, const std::vector<REAL>& local, int Nsystems)
{
  std::vector<REAL> all(Nsystems, 0.0);
  for (size_t i = 0; i < local.size(); i++) {
    size_t g = i * mpi->Nprocs() + mpi->rank();
    if (g < all.size()) all[g] = local[i];
  }
  mpi->allreduce(all);
  return all;
}

std::vector<REAL> Potential::evaluate()
{
  // Potential.cpp:515-551
  if (params.input_file.empty()) PPH_ERROR("Cannot do predictions without checkpoint file (checkpoint_in flag)");
  for (size_t e = 0; e < atom_types.size(); e++) {
    HostNet& n = nets[atom_types[e]];
    PPH_CALL(pp_set_input_scaling(ctx, (int)e, n.params.input_mean.data(), n.params.input_variance.data()));
  }
  PPH_CALL(pp_set_output_scaling(ctx, output_mean, output_variance));
  std::vector<REAL> E(Nlocal, 0.0);
  if (Nlocal) PPH_CALL(pp_evaluate(ctx, E.data()));      // already includes output_mean
  std::vector<REAL> all = gather_in_order(mpi, E, Nsystems);
  if (mpi->io_node()) {
    std::cout << std::endl;
    std::cout << "System         Prediction" << std::endl;
    std::cout << "-----------------------------" << std::endl;
    for (int i = 0; i < Nsystems; i++) std::cout << i + 1 << "            " << all[i] << std::endl;
    std::cout << std::endl;
  }
  return all;
}

// Analysis::KS (Analysis.cpp:149-202): both samples are binned into N = |data1| equal bins over their common range
// (bin width (max-min)/(N-1e-6), Analysis.cpp:108-131), D is the largest gap between the two cumulative
// histograms, and p the alternating Kolmogorov series (501 terms) -- which the reference only evaluates for N > 30.
REAL approximate_KS(const std::vector<REAL>& data1, const std::vector<REAL>& data2, REAL* p_out)
{
  const int N = (int)data1.size();
  *p_out = 0.0;
  if (N == 0 || data2.empty()) return 0.0;
  REAL lo = data1[0], hi = data1[0];
  for (REAL v : data1) { if (v < lo) lo = v; if (v > hi) hi = v; }
  for (REAL v : data2) { if (v < lo) lo = v; if (v > hi) hi = v; }
  const REAL width = (hi - lo) / (N - 1e-6);
  std::vector<int> H1(N, 0), H2(N, 0);
  for (REAL v : data1) H1.at(width > 0 ? (int)((v - lo) / width) : 0)++;
  for (REAL v : data2) H2.at(width > 0 ? (int)((v - lo) / width) : 0)++;
  REAL D = 0, c1 = 0.0, c2 = 0.0;
  for (int i = 0; i < N; i++) {
    c1 += (REAL)H1[i] / data1.size();
    c2 += (REAL)H2[i] / data2.size();
    if (std::fabs(c1 - c2) > D) D = std::fabs(c1 - c2);
  }
  if (D == 0) { *p_out = 1.0; return 0.0; }
  if (N > 30) {
    REAL p = 0.0;
    for (int i = 1; i <= 501; i++) p += std::pow(-1, i - 1) * std::exp(-i * i * (REAL)N * D * D);
    *p_out = 2 * p;
  }
  return D;
}

void Potential::validate()
{
  // Potential.cpp:971-1122, with the approximate Kolmogorov-Smirnov appendix (Analysis.cpp:78-202)
  if (params.input_file.empty()) PPH_ERROR("Cannot do validation without checkpoint file (checkpoint_in flag)");
  std::vector<REAL> E(Nlocal, 0.0);
  for (size_t e = 0; e < atom_types.size(); e++) {
    HostNet& n = nets[atom_types[e]];
    PPH_CALL(pp_set_input_scaling(ctx, (int)e, n.params.input_mean.data(), n.params.input_variance.data()));
  }
  PPH_CALL(pp_set_output_scaling(ctx, output_mean, output_variance));
  if (Nlocal) PPH_CALL(pp_evaluate(ctx, E.data()));
  std::vector<REAL> flag(Nlocal), nat(Nlocal);
  for (int i = 0; i < Nlocal; i++) {
    flag[i] = tags[i] == "train" ? 0 : (tags[i] == "val" ? 1 : (tags[i] == "test" ? 2 : 3));
    nat[i] = natoms[i];
  }
  // with free energies (input_file key or the checkpoints' "FE:" lines) the table shows total energies: the raw
  // network sum and the target both get the per-atom reference energies added back, and output_mean is NOT added
  // (Potential.cpp:981,1001-1018,1046-1051 -> Structure::unravel_Energy)
  const bool unrav = !params.FE.empty();
  std::vector<REAL> fes(Nlocal, 0.0);
  if (unrav)
    for (int i = 0; i < Nlocal; i++) fes[i] = free_energy_sum(params.FE, std::vector<int>(atom_Z[i].begin(), atom_Z[i].end()));
  std::vector<REAL> out = gather_in_order(mpi, E, Nsystems), tg = gather_in_order(mpi, targets, Nsystems);
  std::vector<REAL> fe = gather_in_order(mpi, fes, Nsystems);
  std::vector<REAL> fl = gather_in_order(mpi, flag, Nsystems), na = gather_in_order(mpi, nat, Nsystems);
  if (!mpi->io_node()) return;
  using std::setw;
  std::cout << std::endl;
  std::cout << "System" << setw(20) << "Prediction" << setw(15) << "Target" << setw(18) << "Natom" << setw(15) << "Train" << std::endl;
  std::cout << std::setfill('-') << setw(80) << "-" << std::setfill(' ') << std::endl;
  REAL SSE = 0.0, min_E = 9e9, max_E = 0.0;
  std::cout << std::setprecision(10);
  std::vector<REAL> shown(out), shown_t(tg);
  for (int i = 0; i < Nsystems; i++) {
    REAL Error;
    if (unrav) {
      shown[i] = (out[i] - output_mean) + fe[i];
      shown_t[i] = (targets_shifted ? tg[i] : tg[i] - fe[i]) + fe[i];
      std::cout << i + 1 << "               " << shown[i] << setw(25) << shown_t[i] << setw(15) << (int)na[i] << std::endl;
      Error = shown[i] - shown_t[i];
    } else {
      int f = (int)fl[i];
      const char* t = f == 0 ? "train" : f == 1 ? "val" : f == 2 ? "test" : "other";
      std::cout << setw(4) << i + 1 << setw(22) << out[i] << setw(20) << tg[i] << setw(14) << (int)na[i] << setw(14) << t << std::endl;
      Error = out[i] - tg[i];
    }
    SSE += Error * Error;
    if (std::fabs(Error) < std::fabs(min_E)) min_E = Error;
    if (std::fabs(Error) > std::fabs(max_E)) max_E = Error;
  }
  std::cout << std::setprecision(6);
  std::cout << std::endl;
  std::cout << "Minimum Error =  " << min_E << std::endl;
  std::cout << "Maximum Error =  " << max_E << std::endl;
  std::cout << "RMS Error =  " << std::sqrt(SSE / (double)Nsystems) << std::endl;
  // the reference compares the raw network sums (before output_mean is added back) with the targets -- or, with
  // free energies, the two total-energy columns
  std::vector<REAL> raw(out);
  for (int i = 0; i < Nsystems; i++) raw[i] -= output_mean;
  REAL p = 0.0;
  REAL D = unrav ? approximate_KS(shown, shown_t, &p) : approximate_KS(raw, tg, &p);
  std::cout << std::endl << std::endl;
  std::cout << "-- Approximate Kolmogorov-Smirnov test --" << std::endl;
  std::cout << "  D       =  " << D << std::endl;
  std::cout << "  p value = " << p << std::endl << std::endl;
}

}  // namespace pph
