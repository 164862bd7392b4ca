// pair_style nn on the B200 device path: the reference's LAMMPS plug-in (pair_nn.h:38-89, pair_nn.cpp) with the same
// user-facing commands
//     pair_style nn <Rcut>
//     pair_coeff <type> <checkpoint_file>
// but with PairNN::compute forwarding LAMMPS' own arrays to pp_md_compute instead of copying them into nested
// vectors and looping over atoms on the CPU.
#ifdef PAIR_CLASS
PairStyle(nn, PairNN)
#else
#ifndef LMP_PAIR_NN_B200_H
#define LMP_PAIR_NN_B200_H

#ifdef PPH_LAMMPS_STUB
#include "lammps_stub.h"
#else
#include "pair.h"
#endif
#include <string>
#include <vector>

#include "pph.h"

namespace LAMMPS_NS {

class PairNN : public Pair {
 public:
  explicit PairNN(LAMMPS* lmp);
  virtual ~PairNN();
  virtual void compute(int eflag, int vflag);
  virtual void settings(int narg, char** arg);      // exactly one argument: Rcut (pair_nn.cpp:502-507)
  virtual void coeff(int narg, char** arg);         // type file [cut]       (pair_nn.cpp:226-234)
  void init_style();                                // full list, skin 1.0, rebuilt every 10 steps (pair_nn.cpp:262-285)
  void init_list(int id, NeighList* ptr) { if (id == 1) list = ptr; }
  double init_one(int, int) { return Rcut; }
  // restart record (pair_nn.cpp:322-398): Rcut, then per pair_coeff line its type, the length of the checkpoint file
  // name and the name itself -- reading it replays settings() and coeff()
  void write_restart(FILE* fp);
  void read_restart(FILE* fp);
  void write_restart_settings(FILE*) {}             // nothing, as in the reference (:405-447)
  void read_restart_settings(FILE*) {}
  void write_data(FILE*) {}                         // nothing, as in the reference (:455-474)
  void write_data_all(FILE*) {}
  double single(int, int, int, int, double, double, double, double&) { return 1; }      // pair_nn.cpp:480-485
  double energy_shift() const { return out_mean; }

 private:
  void finish_setup();
  double Rcut, out_mean;
  pp_context* ctx;
  bool ready;
  bool has_free_energy;                       // some pair_coeff file carries an "FE:" line
  std::vector<double> type_free_energy, eatom_tmp;
  std::vector<char> type_has_free_energy;
  std::vector<int> coeff_types;
  std::vector<std::string> coeff_files;       // as given to pair_coeff, for the restart record
  std::vector<pph::HostNet> coeff_nets;
  std::vector<long long> offsets;
  std::vector<int> flat;
};

}  // namespace LAMMPS_NS
#endif
#endif
