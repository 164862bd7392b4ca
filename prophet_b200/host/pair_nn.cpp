// See pair_nn.h.  Reference: pair_nn.cpp:112-188 (compute), 223-259 (coeff), 499-539 (settings),
// Potential::insert_atom_type (Potential.cpp:1214-1253).
#include "pair_nn.h"

#include <cstring>
#include <fstream>

using namespace LAMMPS_NS;

PairNN::PairNN(LAMMPS* lmp) : Pair(lmp), Rcut(0.0), out_mean(0.0), ctx(NULL), ready(false), has_free_energy(false) {}

PairNN::~PairNN()
{
  if (ctx) pp_destroy(ctx);
}

void PairNN::settings(int narg, char** arg)
{
  if (narg != 1) error->all(FLERR, "pair_style nn takes exactly one argument: the cutoff radius");
  Rcut = atof(arg[0]);
  if (!ctx) {
    int device = getenv("LOCAL_RANK") ? atoi(getenv("LOCAL_RANK")) : 0;
    if (pp_create(device, 64, &ctx)) error->all(FLERR, pp_last_error());
  }
}

void PairNN::coeff(int narg, char** arg)
{
  if (narg != 2 && narg != 3) error->all(FLERR, "Format of pair_coeff command is\npair_coeff atom_type parameter_file\n");
  const int this_type = atoi(arg[0]);
  for (size_t i = 0; i < coeff_types.size(); i++)
    if (coeff_types[i] == this_type) return;          // Potential::insert_atom_type: a type is inserted once
  std::ifstream file(arg[1]);
  if (!file.is_open()) error->all(FLERR, (std::string("Could not open parameter file \"") + arg[1] + "\"").c_str());
  pph::HostNet net;
  net.params.read(file);                              // header: rows, element list, means, hidden layers
  net.Z = pph::atomic_number(net.params.current_atom_type);
  net.sizes.clear();
  net.sizes.push_back(net.params.NG());
  for (size_t l = 0; l < net.params.hidden.size(); l++) net.sizes.push_back(net.params.hidden[l]);
  net.sizes.push_back(1);
  net.load(file);
  out_mean = net.params.output_mean;
  if (comm->me == 0) printf("Initializing atom type %d with file %s\n", this_type, arg[1]);
  coeff_types.push_back(this_type);
  coeff_files.push_back(arg[1]);
  coeff_nets.push_back(net);
  ready = false;
  if (coeff_nets.size() == net.params.atomic_symbols.size()) finish_setup();
}

void PairNN::init_style()
{
  // the reference overrides the user's neighbour settings (pair_nn.cpp:268-281): skin 1.0, list rebuilt every 10
  // steps without delay, one FULL list with id 1
  neighbor->cutneighmin = 1.0;
  neighbor->cutneighmax = Rcut;
  neighbor->delay = 0;
  neighbor->every = 10;
  neighbor->skin = 1.0;
  const int irequest = neighbor->request(this, instance_me);
  neighbor->requests[irequest]->pair = 1;
  neighbor->requests[irequest]->id = 1;
  neighbor->requests[irequest]->half = 0;
  neighbor->requests[irequest]->full = 1;
  neighbor->requests[irequest]->occasional = 0;
}

void PairNN::write_restart(FILE* fp)
{
  write_restart_settings(fp);
  if (comm->me != 0) return;
  double rc = Rcut;
  fwrite(&rc, sizeof(double), 1, fp);
  for (size_t i = 0; i < coeff_types.size(); i++) {
    int size = (int)coeff_files[i].size(), type = coeff_types[i];
    fwrite(&type, sizeof(int), 1, fp);
    fwrite(&size, sizeof(int), 1, fp);
    fwrite(coeff_files[i].data(), sizeof(char), (size_t)size, fp);
  }
}

void PairNN::read_restart(FILE* fp)
{
  read_restart_settings(fp);
  // Rank 0 reads, everybody replays (the reference broadcasts each item with MPI_Bcast over `world`; with LAMMPS'
  // MPI that is where the three Bcast calls go -- the stand-in build is single rank)
  double rc = 0.0;
  if (comm->me == 0 && fread(&rc, sizeof(double), 1, fp) != 1) error->all(FLERR, "pair_style nn: truncated restart record");
#ifndef PPH_LAMMPS_STUB
  MPI_Bcast(&rc, 1, MPI_DOUBLE, 0, world);
#endif
  char num[64];
  snprintf(num, sizeof(num), "%.17g", rc);
  char* sarg[1] = {num};
  settings(1, sarg);
  for (int i = 1; i <= atom->ntypes; i++) {
    int type = 0, size = 0;
    if (comm->me == 0 && (fread(&type, sizeof(int), 1, fp) != 1 || fread(&size, sizeof(int), 1, fp) != 1))
      error->all(FLERR, "pair_style nn: truncated restart record");
#ifndef PPH_LAMMPS_STUB
    MPI_Bcast(&type, 1, MPI_INT, 0, world);
    MPI_Bcast(&size, 1, MPI_INT, 0, world);
#endif
    std::string name((size_t)size, '\0');
    if (comm->me == 0 && size > 0 && fread(&name[0], sizeof(char), (size_t)size, fp) != (size_t)size)
      error->all(FLERR, "pair_style nn: truncated restart record");
#ifndef PPH_LAMMPS_STUB
    MPI_Bcast(&name[0], size, MPI_CHAR, 0, world);
#endif
    char tnum[16];
    snprintf(tnum, sizeof(tnum), "%d", type);
    char* carg[2] = {tnum, &name[0]};
    coeff(2, carg);
  }
}

void PairNN::finish_setup()
{
  // Once every element named in the checkpoint header has a pair_coeff line, map LAMMPS types to the header's
  // element order (ascending Z), as Potential::insert_atom_type does (Potential.cpp:1236-1250).
  const pph::FunctionalParams& P = coeff_nets[0].params;
  const int n_elem = (int)P.atomic_symbols.size();
  std::vector<double> g1, g2, g3, g4;
  for (size_t r = 0; r < P.G1.size(); r++) g1.insert(g1.end(), P.G1[r].begin(), P.G1[r].end());
  for (size_t r = 0; r < P.G2.size(); r++) g2.insert(g2.end(), P.G2[r].begin(), P.G2[r].end());
  for (size_t r = 0; r < P.G3.size(); r++) g3.insert(g3.end(), P.G3[r].begin(), P.G3[r].end());
  for (size_t r = 0; r < P.G4.size(); r++) g4.insert(g4.end(), P.G4[r].begin(), P.G4[r].end());
  const double rc = Rcut > 0 ? Rcut : P.Rcut;
  if (pp_set_descriptors(ctx, n_elem, (int)P.G1.size(), g1.data(), (int)P.G2.size(), g2.data(), (int)P.G3.size(), g3.data(),
                         (int)P.G4.size(), g4.data(), rc))
    error->all(FLERR, pp_last_error());
  int max_type = 0;
  for (size_t i = 0; i < coeff_types.size(); i++) max_type = std::max(max_type, coeff_types[i]);
  std::vector<int> type_to_elem(max_type + 1, 0);
  for (size_t i = 0; i < coeff_nets.size(); i++) {
    pph::HostNet& n = coeff_nets[i];
    int e = -1;
    for (int k = 0; k < n_elem; k++)
      if (P.atomic_symbols[k] == n.params.current_atom_type) e = k;
    if (e < 0) error->all(FLERR, "pair_coeff file's element is not in the potential's element list");
    type_to_elem[coeff_types[i]] = e;
    std::vector<int> tf;
    for (size_t k = 0; k < n.node_transfer.size(); k++) tf.push_back(pph::transfer_id(n.node_transfer[k]));
    if (pp_set_network(ctx, e, (int)n.sizes.size() - 1, n.sizes.data(), tf.data(), n.weights.data(), n.params.input_mean.data(),
                       n.params.input_variance.data()))
      error->all(FLERR, pp_last_error());
  }
  if (pp_set_output_scaling(ctx, coeff_nets[0].params.output_mean, coeff_nets[0].params.output_variance)) error->all(FLERR, pp_last_error());
  if (pp_md_set_types(ctx, max_type, type_to_elem.data())) error->all(FLERR, pp_last_error());
  // free energies: each pair_coeff file contributes its own element's "FE:" entry (Functional_params.cpp:257-264);
  // lammps type -> element symbol as pair_nn.cpp:246 records it
  has_free_energy = false;
  type_free_energy.assign(max_type + 1, 0.0);
  type_has_free_energy.assign(max_type + 1, 0);
  for (size_t i = 0; i < coeff_nets.size(); i++) {
    const pph::FunctionalParams& fp = coeff_nets[i].params;
    std::map<std::string, pph::REAL>::const_iterator it = fp.FE.find(fp.current_atom_type);
    if (it != fp.FE.end()) {
      has_free_energy = true;
      type_free_energy[coeff_types[i]] = it->second;
      type_has_free_energy[coeff_types[i]] = 1;
    }
  }
  ready = true;
}

void PairNN::compute(int eflag, int vflag)
{
  if (eflag || vflag) ev_setup(eflag, vflag);
  else evflag = vflag_fdotr = 0;
  if (!ready) error->all(FLERR, "pair_style nn: not every element of the potential has a pair_coeff line");
  const int nlocal = atom->nlocal;
  // LAMMPS' paged full neighbour list -> one CSR list.  Like the reference this assumes ilist[ii] == ii
  // (pair_nn.cpp:151-157) and unmasked neighbour indices.
  offsets.resize(nlocal + 1);
  offsets[0] = 0;
  for (int ii = 0; ii < list->inum; ii++) offsets[ii + 1] = offsets[ii] + list->numneigh[ii];
  flat.resize((size_t)offsets[list->inum] + 1);
  for (int ii = 0; ii < list->inum; ii++)
    memcpy(flat.data() + offsets[ii], list->firstneigh[ii], sizeof(int) * (size_t)list->numneigh[ii]);
  double etotal = 0.0;
  if (has_free_energy) {
    // Checkpoints with "FE:" lines (Potential::evaluate_MD, Potential.cpp:1159-1162 -> Structure::unravel_Energy,
    // Structure.cpp:152-199): every atom's energy becomes (E_i + S) / N, with S the free energies summed over ALL
    // atoms this rank sees (locals and ghosts) and N their number.  Odd, but it is what the reference returns.
    const int ntotal = nlocal + atom->nghost;
    double S = 0.0;
    for (int k = 0; k < ntotal; k++) {
      const int t = atom->type[k];
      if (t < 0 || t >= (int)type_free_energy.size() || !type_has_free_energy[t])
        error->all(FLERR, "No Free Energy supplied for an atom type present in the system");
      S += type_free_energy[t];
    }
    eatom_tmp.assign(nlocal, 0.0);
    if (pp_md_compute(ctx, atom->x[0], atom->type, nlocal, atom->nghost, offsets.data(), flat.data(), atom->f[0],
                      eatom_tmp.data(), &etotal, NULL, NULL))
      error->all(FLERR, pp_last_error());
    for (int ii = 0; ii < nlocal; ii++) {
      const double Ei = (eatom_tmp[ii] + S) / ntotal;
      if (eflag_global) eng_vdwl += Ei;
      if (eflag_atom) eatom[ii] += Ei;
    }
    if (eflag_global && comm->me == 0) eng_vdwl += out_mean;
    if (vflag_fdotr) virial_fdotr_compute();
    return;
  }
  if (pp_md_compute(ctx, atom->x[0], atom->type, nlocal, atom->nghost, offsets.data(), flat.data(), atom->f[0],
                    eflag_atom ? eatom : NULL, &etotal, NULL, NULL))
    error->all(FLERR, pp_last_error());
  if (eflag_global) {
    // per-atom energies summed on every rank, the shift added once on rank 0 (pair_nn.cpp:156-179)
    eng_vdwl += etotal - out_mean;
    if (comm->me == 0) eng_vdwl += out_mean;
  }
  if (vflag_fdotr) virial_fdotr_compute();
}
