// Minimal stand-in for the LAMMPS classes pair_nn touches, used ONLY to compile-check and unit-test the pair style
// without a LAMMPS tree (LAMMPS is not installed in this image).  The member names and meanings follow the public
// LAMMPS Pair / Atom / NeighList interfaces the reference's pair_nn.cpp uses (pair_nn.cpp:112-188, 223-285, 499-539).
// Building against a real LAMMPS: drop this header and include LAMMPS' own pair.h / atom.h / neigh_list.h instead.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <string>

namespace LAMMPS_NS {

struct Atom {
  double** x;
  double** f;
  int* type;
  int nlocal, nghost, ntypes;
};

struct NeighList {
  int inum;
  int* ilist;
  int* numneigh;
  int** firstneigh;
};

struct Comm { int me; };

// neighbour-list request as pair_nn.cpp:262-285 fills it in (pre-2020 LAMMPS API)
struct NeighRequest { int pair, id, half, full, occasional; };
struct Neighbor {
  double cutneighmin, cutneighmax, skin;
  int delay, every;
  NeighRequest* requests[4];
  NeighRequest storage[4];
  int nrequest;
  Neighbor() : cutneighmin(0), cutneighmax(0), skin(0), delay(0), every(0), nrequest(0)
  {
    for (int i = 0; i < 4; i++) requests[i] = &storage[i];
  }
  int request(void*, int) { return nrequest < 3 ? nrequest++ : 3; }
};

struct Error {
  void all(const char* file, int line, const char* msg)
  {
    fprintf(stderr, "ERROR: %s (%s:%d)\n", msg, file, line);
    exit(1);
  }
};

struct LAMMPS {
  Atom* atom;
  Comm* comm;
  Error* error;
  Neighbor* neighbor;
};

class Pair {
 public:
  explicit Pair(LAMMPS* lmp) : lmp_(lmp), atom(lmp->atom), comm(lmp->comm), error(lmp->error), neighbor(lmp->neighbor), list(NULL)
  {
    instance_me = 0;
    eng_vdwl = 0.0;
    eatom = NULL;
    eflag_global = eflag_atom = vflag_fdotr = evflag = 0;
    allocated = 0;
    setflag = NULL;
    cutsq = NULL;
  }
  virtual ~Pair() {}
  virtual void compute(int, int) = 0;
  virtual void settings(int, char**) = 0;
  virtual void coeff(int, char**) = 0;
  double eng_vdwl;
  double* eatom;
  int eflag_global, eflag_atom, vflag_fdotr, evflag;

 protected:
  void ev_setup(int eflag, int vflag)
  {
    evflag = 1;
    eflag_global = eflag & 1;
    eflag_atom = eflag & 2;
    vflag_fdotr = vflag & 2 ? 1 : 0;
    if (eflag_global) eng_vdwl = 0.0;
    if (eflag_atom && eatom)
      for (int i = 0; i < atom->nlocal; i++) eatom[i] = 0.0;
  }
  void virial_fdotr_compute() {}
  LAMMPS* lmp_;
  Atom* atom;
  Comm* comm;
  Error* error;
  Neighbor* neighbor;
  NeighList* list;
  int instance_me;
  int allocated;
  int** setflag;
  double** cutsq;
};

}  // namespace LAMMPS_NS
#define FLERR __FILE__, __LINE__
