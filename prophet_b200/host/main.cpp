// PROPhet_b200: the reference's command line (main.cpp:20-98) on the B200 device path.
//   PROPhet_b200 [-in input_file] [-train | -run | -validate | -optimize]
// Multi-GPU: start one process per GPU with RANK / WORLD_SIZE / LOCAL_RANK set (torchrun's variables) or
// PP_RANK / PP_WORLD_SIZE / PP_LOCAL_RANK; structures are sharded round robin as Setup.cpp:582 does over MPI ranks.
#include <cstring>

#include "pph.h"

int main(int argc, char* argv[])
{
  std::string input_file = "input_file";
  std::string run_type = "train";
  for (int i = 1; i < argc; i++) {
    if (!strcmp(argv[i], "-in") || !strcmp(argv[i], "-input")) {
      if (argc > i + 1) input_file = argv[++i];
    } else if (!strcmp(argv[i], "-run")) run_type = "run";
    else if (!strcmp(argv[i], "-train")) run_type = "train";
    else if (!strcmp(argv[i], "-validate")) run_type = "validate";
    else if (!strcmp(argv[i], "-optimize")) run_type = "optimize";
    else if (!strcmp(argv[i], "-init")) run_type = "initialize";      // accepted; for a potential the reference then
    else if (!strcmp(argv[i], "-forces")) run_type = "forces";        // only sets everything up and exits (main.cpp:32,43,66-77)
    else PPH_ERROR(std::string("Command line argument \"") + argv[i] + "\" not recognized");
  }
  pph::Parallel mpi;
  pph::Setup run_params(input_file, &mpi);
  pph::Potential pot(run_params, &mpi);
  if (run_type == "train") pot.train();
  else if (run_type == "run") pot.evaluate();
  else if (run_type == "validate") pot.validate();
  else if (run_type == "optimize") pot.optimize_Gs();
  return 0;
}
