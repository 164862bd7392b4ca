// Host-side C++ layer above the C ABI (include/prophet_b200.h): the classes a PROPhet user knows -- Setup,
// Functional_params, System/Structure input, Potential, Optimizer, Parallel -- re-implemented so that the
// reference's input_file grammar, checkpoint format, CLI flags and stdout tables keep working while all of the
// arithmetic of the hot path runs on the GPU.  Nothing here computes descriptors, networks or forces on the CPU.
//
// Mirrors (file:line in the reference's src/):
//   Setup            Setup.cpp:50-724         Functional_params  Functional_params.cpp:52-372
//   System/readers   System.cpp:51-160, CUSTOM.cpp:81-160, VASP.cpp:78-176
//   Potential        Potential.cpp:60-693, 971-1122            Optimizer  Optimizer.cpp:193-593, 884-1064
//   Parallel         Parallel.h:93-262
#pragma once
#include <deque>
#include <iostream>
#include <map>
#include <string>
#include <vector>

#include "../../include/prophet_b200.h"

namespace pph {

typedef double REAL;

// Error.cpp:49-81: print where and why, then stop every rank.
[[noreturn]] void fatal(const std::string& msg, const char* file, int line);
#define PPH_ERROR(msg) ::pph::fatal((msg), __FILE__, __LINE__)
#define PPH_CALL(call)                                                        \
  do {                                                                        \
    if ((call) != 0) PPH_ERROR(std::string("prophet_b200: ") + pp_last_error()); \
  } while (0)

// ---- Atom.cpp:168-185 ----------------------------------------------------------------------------------------
int atomic_number(const std::string& symbol);     // 0 if unknown
std::string atomic_symbol(int Z);

// ---- Functional_params ---------------------------------------------------------------------------------------
struct FunctionalParams {
  FunctionalParams();
  // defaults: Functional_params.cpp:56-103
  std::string input_file, output_file, network_type, transfer_function, training_algorithm, output;
  std::vector<std::string> inputs;
  std::vector<int> hidden;
  int Niterations, Ncheckpoint, Nprint, Ncycles, Nradial, Nangular, Nbackup;
  bool Tbackup, SGD, output_precondition, input_precondition, print_gvectors, is_MD;
  REAL Rcut, threshold, regularization, sd_momentum, line_min_epsilon, lambda_max, debug, dropout;
  REAL T0, Tf, dT, step_size;                // -optimize: annealing schedule and Gaussian step (Functional_params.cpp:78-82)
  int Nsteps_per_T;
  std::vector<std::vector<REAL> > G1, G2, G3, G4;
  std::map<std::string, REAL> FE;
  // per-element state carried by a checkpoint
  std::string current_atom_type;
  std::vector<std::string> atomic_symbols;
  std::vector<int> atomic_numbers;
  std::vector<REAL> input_mean, input_variance;
  REAL output_mean, output_variance;

  int NG() const;
  void auto_rows();                          // Structure::init_G default rows (Structure.cpp:818-851)
  void print(std::ostream& out) const;       // checkpoint header, Functional_params.cpp:133-193
  void read(std::istream& in);               // Functional_params.cpp:207-372
};

// ---- structures as the readers deliver them ------------------------------------------------------------------
struct StructureData {
  double cell[9];                    // rows a, b, c
  std::vector<double> pos;           // [n][3] fractional AFTER Set_Convert/Normalize (what Structure::pos holds)
  std::vector<int> Z;
  double target;
  std::string train;                 // "train" | "val" | "test" | other
  int natoms() const { return (int)Z.size(); }
};

struct SystemSpec { std::map<std::string, std::string> files; };

// ---- Setup ----------------------------------------------------------------------------------------------------
class Parallel;
class Setup {
 public:
  Setup(const std::string& input_file, Parallel* mpi, bool quiet = false);
  FunctionalParams F;
  int Nsystems() const { return (int)systems.size(); }
  const SystemSpec& system(int i) const { return systems.at(i); }
  bool is_potential() const { return my_is_potential; }
  int Ntotal_systems;
  StructureData load(int i) const;   // System::System (System.cpp:51-150): reader by `code`, target, train tag

 private:
  void read_input(const std::string& filename);
  void split_systems();
  static std::string clean_line(std::string line);
  std::vector<SystemSpec> systems;
  std::map<std::string, std::string> property_map;
  Parallel* mpi;
  bool my_is_potential;
};

// sum of the `free_energy` entries over a structure's atoms (Structure::train_Local / unravel_Energy, Structure.cpp:110-199)
REAL free_energy_sum(const std::map<std::string, REAL>& FE, const std::vector<int>& Z);
StructureData read_prophet_xml(int index);                    // CUSTOM.cpp:81-160 ("PROPhet.xml" in the CWD)
StructureData read_vasprun(const std::string& dir);           // VASP.cpp:78-176

// ---- Parallel -------------------------------------------------------------------------------------------------
// Rank/size come from the launcher's environment (RANK / WORLD_SIZE / LOCAL_RANK as torchrun sets them, or
// PP_RANK / PP_WORLD_SIZE); rank 0 publishes the ncclUniqueId in a rendezvous file.  The collectives themselves
// are NCCL, through the C ABI.
class Parallel {
 public:
  Parallel();
  int rank() const { return my_rank; }
  int Nprocs() const { return my_nprocs; }
  int local_rank() const { return my_local_rank; }
  bool io_node() const { return my_rank == 0; }
  void attach(pp_context* ctx);                 // pp_comm_init on this context
  // Reduce(SUM) + Bcast of small host vectors (Potential.cpp:73-74,170-172 ...); identity before attach / for one rank
  void allreduce(std::vector<REAL>& v);
  REAL allreduce(REAL v);

 private:
  int my_rank, my_nprocs, my_local_rank;
  pp_context* ctx;
};

// ---- one element's network on the host -------------------------------------------------------------------------
struct HostNet {
  int Z;
  FunctionalParams params;                 // header state of this element (means, rows, hidden ...)
  std::vector<int> sizes;                  // NG, hidden..., 1
  std::vector<std::string> node_transfer;  // per node, output node "linear"
  std::vector<REAL> weights;               // flat, layer-major, node-major, [w..., b]
  bool preconditioned;                     // loaded from a checkpoint (NN.h:100-103)
  int Nparameters() const { return (int)weights.size(); }
  void init_random();                      // NN_node.cpp:52-62
  void load(std::istream& in);             // NN.cpp:158-200
  void print(std::ostream& out) const;     // NN.cpp:828-847 + NN_node.cpp:232-243
};
int transfer_id(const std::string& name);

// ---- Optimizer ------------------------------------------------------------------------------------------------
class Optimizer {
 public:
  Optimizer(Parallel* mpi, const FunctionalParams& params, int Nsystems);
  void init(const std::vector<REAL>& x0);                 // starting point + table header (Optimizer.cpp:242-289)
  // one call per epoch (Optimizer.cpp:884-1064).  Error/grad are already summed over ranks (the sum inside pp_epoch).
  void update_network(REAL Error, const std::vector<REAL>& grad);
  void set_val_sse(REAL sse, long long n) { val_sse = sse; val_count = n; }
  void set_early_stop(bool e) { report_validation = e; }
  bool is_converged() const { return done; }
  bool is_checkpoint() const { return want_checkpoint; }
  bool is_backup() const { return want_backup; }
  int iteration() const { return epoch; }
  REAL get_error() const { return *errors.newest(0); }
  const std::vector<REAL>& parameters() const { return current; }
  REAL alpha() const { return pt.alpha; }          // perturbation amplitude (Optimizer.h:130-137); unused by Potential::train
  REAL dphi_dt() const { return pt.dphi; }
  bool do_print;

 private:
  enum Method { kRprop = 0, kDescent = 1, kDescentLine = 2, kLbfgs = 3 };
  // fixed-capacity FIFO of equal-length rows in one flat allocation: the kept part of the trajectory
  struct Ring {
    std::vector<REAL> buf;
    int w, cap, first, n;
    Ring() : w(0), cap(0), first(0), n(0) {}
    void reset(int width, int capacity);
    int size() const { return n; }
    REAL* append();
    void drop_oldest();
    const REAL* oldest(int i) const;
    const REAL* newest(int age) const;
    REAL* newest_mut();
  };
  struct Pair { std::vector<REAL> s, y; REAL rho; };        // one L-BFGS correction pair
  struct LineSearch {                                       // state of the line minimisation (see line_search)
    enum Phase { kFresh, kExpanding, kBracketed } phase;
    int npts;
    REAL a[2], s[2];                   // distances along the line and directional slopes of the kept points
    REAL ok_E, ok_a, ok_s;             // last accepted point while expanding
    REAL c_a, c_s, d_a;                // previous best point and the one before it (interpolation / safeguards)
    bool bisected;
    REAL resolution;
  };
  struct Perturbation {                                     // lambda_max != 0 (Optimizer.cpp:915-938, 1008-1043)
    REAL ceiling, amplitude_now, alpha, dphi, phase, rate, best_E, best_norm, floor_E;
    bool enabled;
    int cycles, t_on;
    std::vector<REAL> best_x;
  };
  std::vector<REAL> step_rprop();
  std::vector<REAL> step_descent();
  std::vector<REAL> step_descent_line();
  std::vector<REAL> step_lbfgs();
  std::vector<REAL> dispatch();
  std::vector<REAL> enter(Method m);
  std::vector<REAL> along_line(REAL t) const;
  void start_line_at_current(const REAL* direction_source);
  REAL line_search(REAL tolerance);
  void perturbation_tick(REAL shown_E);

  Parallel* mpi;
  FunctionalParams F;
  Method method;
  bool may_switch, report_validation, done, want_checkpoint, want_backup, line_done;
  int n_structures, n_par, epoch;
  REAL val_sse;
  long long val_count;
  Ring points, slopes, errors;
  std::vector<REAL> current;
  // Rprop
  REAL first_size, size_ceiling, size_floor, restart_size;
  std::vector<REAL> rp_size, rp_move;
  std::vector<signed char> rp_live;
  bool rp_started;
  // descent / line methods
  REAL step_len;
  std::vector<REAL> dir, line_origin, line_origin_slope;
  bool dir_set;
  int lbfgs_memory;
  std::vector<Pair> pairs;
  LineSearch ls;
  Perturbation pt;
};

REAL approximate_KS(const std::vector<REAL>& data1, const std::vector<REAL>& data2, REAL* p_out);   // Analysis.cpp:149-202

// ---- Potential ------------------------------------------------------------------------------------------------
class Potential {
 public:
  Potential(const Setup& setup, Parallel* mpi);
  ~Potential();
  REAL train();                       // Potential.cpp:607-693
  std::vector<REAL> evaluate();       // :515-551
  void validate();                    // :971-1122
  void optimize_Gs();                 // :709-877  Monte-Carlo search over the descriptor row parameters
  void save();                        // :296-317
  void backup(int niter);             // :323-345
  pp_context* context() { return ctx; }
  int Ntotal_params;

 private:
  void Precondition();                // :387-502
  void upload_networks();
  void upload_descriptor_rows();      // pp_set_descriptors from params.G1..G4
  std::vector<REAL> all_parameters() const;
  void set_all_parameters(const std::vector<REAL>& x);
  Parallel* mpi;
  FunctionalParams params;
  pp_context* ctx;
  std::vector<int> atom_types;                 // ascending Z
  std::map<int, HostNet> nets;
  std::vector<REAL> targets;                   // local systems
  std::vector<std::string> tags;
  std::vector<int> natoms;
  std::vector<std::vector<unsigned char> > atom_Z;   // per local system, per atom: atomic number (for the free-energy sums)
  bool targets_shifted;                        // `free_energy` was in the input_file: targets had the sums taken off at load
  int Nsystems, Nlocal, Ntrain, Nval, Nother;
  bool early_stop;
  REAL output_mean, output_variance;
};

}  // namespace pph
