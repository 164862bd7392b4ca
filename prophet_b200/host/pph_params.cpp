// Functional_params, Setup and the structure readers of the host layer (see pph.h for the reference map).
#include "pph.h"

#include <sys/stat.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <fstream>
#include <iomanip>
#include <limits>
#include <sstream>

namespace pph {

void fatal(const std::string& msg, const char* file, int line)
{
  // Error.cpp:49-81: location, message, then stop
  std::cerr << std::endl << "ERROR in " << file << " at line " << line << ":" << std::endl << "   " << msg << std::endl << std::endl;
  std::cerr.flush();
  std::cout.flush();
  exit(1);
}

// ---- Atom table (Atom.cpp:168-185) -----------------------------------------------------------------------------
static const char* kSymbols[] = {
    "X",  "H",  "He", "Li", "Be", "B",  "C",  "N",  "O",  "F",  "Ne", "Na", "Mg", "Al", "Si", "P",  "S",  "Cl", "Ar", "K",
    "Ca", "Sc", "Ti", "V",  "Cr", "Mn", "Fe", "Co", "Ni", "Cu", "Zn", "Ga", "Ge", "As", "Se", "Br", "Kr", "Rb", "Sr", "Y",
    "Zr", "Nb", "Mo", "Tc", "Ru", "Rh", "Pd", "Ag", "Cd", "In", "Sn", "Sb", "Te", "I",  "Xe", "Cs", "Ba", "La", "Ce", "Pr",
    "Nd", "Pm", "Sm", "Eu", "Gd", "Tb", "Dy", "Ho", "Er", "Tm", "Yb", "Lu", "Hf", "Ta", "W",  "Re", "Os", "Ir", "Pt", "Au",
    "Hg", "Tl", "Pb", "Bi", "Po", "At", "Rn", "Fr", "Ra", "Ac", "Th", "Pa", "U",  "Np", "Pu", "Am", "Cm", "Bk", "Cf", "Es",
    "Fm", "Md", "No", "Lr", "Rf", "Db", "Sg", "Bh", "Hs", "Mt", "Ds", "Rg", "Cn", "Nh", "Fl", "Mc", "Lv", "Ts", "Og"};
static const int kNsymbols = (int)(sizeof(kSymbols) / sizeof(kSymbols[0]));

int atomic_number(const std::string& symbol_in)
{
  std::string s;
  for (size_t i = 0; i < symbol_in.size(); i++)
    if (!isspace((unsigned char)symbol_in[i])) s.push_back(symbol_in[i]);
  if (s.empty()) return 0;
  s[0] = (char)toupper((unsigned char)s[0]);
  for (size_t i = 1; i < s.size(); i++) s[i] = (char)tolower((unsigned char)s[i]);
  for (int z = 1; z < kNsymbols; z++)
    if (s == kSymbols[z]) return z;
  return 0;
}

std::string atomic_symbol(int Z) { return (Z > 0 && Z < kNsymbols) ? kSymbols[Z] : "X"; }

// ---- Functional_params ---------------------------------------------------------------------------------------
FunctionalParams::FunctionalParams()
{
  input_file = "";
  output_file = "functional.save";
  network_type = "neural_network";
  transfer_function = "tanh_spline";
  training_algorithm = "default";
  Niterations = std::numeric_limits<int>::max();
  Ncheckpoint = 100;
  Nprint = 100;
  Ncycles = 1;
  lambda_max = 0;
  line_min_epsilon = 0;
  sd_momentum = 0.0;
  SGD = false;
  Rcut = 6.0;
  Nradial = 8;
  Nangular = 8;
  is_MD = false;
  regularization = 0.0;
  output_precondition = false;
  input_precondition = true;
  output_mean = 0;
  output_variance = 1;
  Nbackup = 0;
  Tbackup = false;
  dropout = 1.0;
  T0 = 1; Tf = 0.1; dT = 0.98; Nsteps_per_T = 50; step_size = 0.2;
  print_gvectors = false;
  threshold = 0.0;
  debug = 0.0;
}

int FunctionalParams::NG() const
{
  const int nt = (int)atomic_numbers.size();
  return nt * (int)(G1.size() + G2.size()) + nt * (nt + 1) / 2 * (int)(G3.size() + G4.size());
}

void FunctionalParams::auto_rows()
{
  // Structure::init_G, Structure.cpp:818-851: generated only when no row of any kind was given
  if (G1.size() + G2.size() + G3.size() + G4.size() != 0) return;
  if (Nradial) {
    REAL stepsize = Rcut / (Nradial - 1);
    for (double rs = 0.0; rs <= Rcut; rs += stepsize) {
      std::vector<REAL> t;
      t.push_back(Rcut);
      t.push_back(0);
      t.push_back(11.5129254649702 / (4 * stepsize * stepsize));
      t.push_back(rs);
      G2.push_back(t);
    }
  }
  if (Nangular) {
    double stepsize = 2 * 15.0 / (Nangular - 2.0);
    for (double xi = 1.0; xi <= 16.001; xi += stepsize) {
      std::vector<REAL> t;
      t.push_back(Rcut);
      t.push_back(0);
      t.push_back(4.60517018598809 / (Rcut * Rcut));
      t.push_back(xi);
      t.push_back(1);
      G4.push_back(t);
      t.back() = -1;
      G4.push_back(t);
    }
  }
}

void FunctionalParams::print(std::ostream& out) const
{
  // statement-for-statement the stream manipulations of Functional_params.cpp:133-193, so that the sticky
  // setprecision(12) reaches the means / scales lines exactly as in the reference's files
  out << network_type << std::endl;
  for (size_t i = 0; i < inputs.size(); i++) out << inputs[i] << "  ";
  out << std::endl;
  if (!current_atom_type.empty()) {
    out << current_atom_type << ":  ";
    for (size_t i = 0; i < atomic_symbols.size(); i++) out << atomic_symbols[i] << " ";
    out << std::endl;
    if (!FE.empty()) {
      out.precision(9);
      std::map<std::string, REAL>::const_iterator it = FE.find(current_atom_type);
      out << "FE:  " << (it == FE.end() ? 0.0 : it->second) << std::endl;
      out.precision(6);
    }
    out << G1.size() + G2.size() + G3.size() + G4.size() << std::endl;
    // (the reference indexes a third element of the 2-element G1 rows here and aborts; we write the two that exist)
    for (size_t i = 0; i < G1.size(); i++) out << std::setprecision(12) << "G1 " << G1[i][0] << " " << G1[i][1] << std::endl;
    for (size_t i = 0; i < G2.size(); i++)
      out << std::setprecision(12) << "G2 " << G2[i][0] << " " << G2[i][1] << " " << G2[i][2] << " " << G2[i][3] << std::endl;
    for (size_t i = 0; i < G3.size(); i++)
      out << std::setprecision(12) << "G3 " << G3[i][0] << " " << G3[i][1] << " " << G3[i][2] << " " << G3[i][3] << " " << G3[i][4] << std::endl;
    for (size_t i = 0; i < G4.size(); i++)
      out << std::setprecision(12) << "G4 " << G4[i][0] << " " << G4[i][1] << " " << G4[i][2] << " " << G4[i][3] << " " << G4[i][4] << std::endl;
  }
  for (size_t i = 0; i < input_mean.size(); i++) out << input_mean[i] << "  ";
  out << std::endl;
  for (size_t i = 0; i < input_variance.size(); i++) out << input_variance[i] << "  ";
  out << std::endl;
  out << output << std::endl;
  out << output_mean << std::endl;
  out << output_variance << std::endl;
  if (hidden.size()) {
    for (size_t i = 0; i < hidden.size(); i++) out << hidden[i] << "  ";
  } else {
    out << 0;
  }
  out << std::endl;
}

void FunctionalParams::read(std::istream& in)
{
  std::string line, input;
  std::istringstream Line;
  bool is_potential = false;
  REAL values;
  std::getline(in, line);
  network_type = line;
  std::getline(in, line);
  Line.str(line);
  std::vector<std::string> names;
  Line >> input;
  if (input.empty()) PPH_ERROR("Could not read network inputs from file");
  names.push_back(input);
  if (input == "structure") is_potential = true;
  while (Line >> input) {
    names.push_back(input);
    if (input == "structure") is_potential = true;
  }
  inputs = names;
  Line.clear();
  if (is_potential) {
    atomic_symbols.clear();
    atomic_numbers.clear();
    std::getline(in, line);
    Line.str(line);
    std::string symbol;
    Line >> symbol;
    if (!symbol.empty()) symbol.erase(symbol.end() - 1, symbol.end());
    current_atom_type = symbol;
    while (Line >> symbol) {
      int z = atomic_number(symbol);
      atomic_symbols.push_back(atomic_symbol(z));
      atomic_numbers.push_back(z);
    }
    Line.clear();
    std::getline(in, line);
    if (line.find("FE") != std::string::npos) {
      Line.str(line);
      REAL fe;
      std::string tmp;
      Line >> tmp >> fe;
      FE[current_atom_type] = fe;
      Line.clear();
      std::getline(in, line);
    }
    Line.str(line);
    int Nfunctions = 0, count = 1;
    Line >> Nfunctions;
    G1.clear(); G2.clear(); G3.clear(); G4.clear();
    Line.clear();
    REAL Rcut_max = 0;
    std::string type;
    while (count <= Nfunctions) {
      std::getline(in, line);
      Line.str(line);
      std::vector<REAL> t;
      Line >> type;
      while (Line >> values) t.push_back(values);
      Line.clear();
      if (type == "G1") G1.push_back(t);
      else if (type == "G2") G2.push_back(t);
      else if (type == "G3") G3.push_back(t);
      else if (type == "G4") G4.push_back(t);
      else PPH_ERROR("Symmetry functions not read correctly from checkpoint file");
      count++;
      if (!t.empty() && t[0] > Rcut_max) Rcut_max = t[0];
    }
    if (Rcut_max > 0) Rcut = Rcut_max;
  }
  input_mean.clear();
  input_variance.clear();
  const std::string numbers = " \t-+0123456789eEdD.";
  std::getline(in, line);
  if (line.find_first_not_of(numbers) != std::string::npos) PPH_ERROR("Input means read incorrectly");
  Line.str(line);
  while (Line >> values) input_mean.push_back(values);
  Line.clear();
  std::getline(in, line);
  if (line.find_first_not_of(numbers) != std::string::npos) PPH_ERROR("Input variances read incorrectly");
  Line.str(line);
  while (Line >> values) input_variance.push_back(values);
  Line.clear();
  if (input_mean.size() != input_variance.size() || input_mean.empty()) PPH_ERROR("Input means/variances read incorrectly");
  std::getline(in, line);
  Line.str(line);
  Line >> output;
  Line.clear();
  std::getline(in, line);
  Line.str(line);
  Line >> output_mean;
  Line.clear();
  std::getline(in, line);
  Line.str(line);
  Line >> output_variance;
  Line.clear();
  std::getline(in, line);
  Line.str(line);
  if (line.empty()) PPH_ERROR("Could not read node configuration from save file");
  int h;
  hidden.clear();
  while (Line >> h) hidden.push_back(h);
  if (hidden.size() == 1 && hidden[0] == 0) hidden.clear();
  Line.clear();
}

// ---- readers --------------------------------------------------------------------------------------------------
static std::string inner_last(const std::string& block, const std::string& tag)
{
  // text of the LAST <tag ...>text</tag> in block (the reference takes .back() of the matches)
  const std::string open = "<" + tag, close = "</" + tag + ">";
  size_t pos = std::string::npos, from = 0;
  for (;;) {
    size_t p = block.find(open, from);
    if (p == std::string::npos) break;
    char nx = p + open.size() < block.size() ? block[p + open.size()] : '>';
    if (nx == '>' || nx == ' ') pos = p;
    from = p + 1;
  }
  if (pos == std::string::npos) return "";
  size_t gt = block.find('>', pos);
  size_t end = block.find(close, gt);
  if (gt == std::string::npos || end == std::string::npos) return "";
  return block.substr(gt + 1, end - gt - 1);
}

static void parse3(const std::string& text, double* v)
{
  std::istringstream s(text);
  for (int k = 0; k < 3; k++)
    if (!(s >> v[k])) PPH_ERROR("could not read three numbers from '" + text + "'");
}

static void finish_cartesian(StructureData& sd, const std::vector<double>& cart)
{
  // Structure::Set_Convert + Normalize (Structure.cpp:1159-1279) through the C ABI (bit-identical arithmetic)
  sd.pos.resize(cart.size());
  PPH_CALL(pp_cart_to_frac(sd.cell, cart.data(), (int)(cart.size() / 3), sd.pos.data()));
}

StructureData read_prophet_xml(int index)
{
  // The reference rescans PROPhet.xml from the top for every system (CUSTOM.cpp:95-105); the file is indexed once here.
  static std::vector<std::string> blocks;
  static bool loaded = false;
  if (!loaded) {
    std::ifstream f("PROPhet.xml");
    if (!f.is_open()) PPH_ERROR("Could not open PROPhet.xml");
    std::string line, cur;
    bool in = false;
    while (std::getline(f, line)) {
      if (!in && line.find("<system id=\"") != std::string::npos) { in = true; cur.clear(); }
      if (in) {
        cur += line;
        cur += "\n";
        if (line.find("</system>") != std::string::npos) { blocks.push_back(cur); in = false; }
      }
    }
    loaded = true;
  }
  if (index < 0 || index >= (int)blocks.size()) PPH_ERROR("system index out of range in PROPhet.xml");
  const std::string& b = blocks[index];
  StructureData sd;
  std::string tr = inner_last(b, "train");
  // trim
  size_t s0 = tr.find_first_not_of(" \t\n"), s1 = tr.find_last_not_of(" \t\n");
  sd.train = (s0 == std::string::npos) ? "train" : tr.substr(s0, s1 - s0 + 1);
  std::string lat = inner_last(b, "lattice");
  parse3(inner_last(lat, "a"), sd.cell);
  parse3(inner_last(lat, "b"), sd.cell + 3);
  parse3(inner_last(lat, "c"), sd.cell + 6);
  std::string atoms = inner_last(b, "atoms");
  std::vector<double> cart;
  size_t from = 0;
  for (;;) {
    size_t p = atoms.find("<atom ", from);
    if (p == std::string::npos) break;
    size_t q0 = atoms.find("specie=\"", p);
    size_t q1 = atoms.find('"', q0 + 8);
    std::string sym = atoms.substr(q0 + 8, q1 - q0 - 8);
    size_t gt = atoms.find('>', q1), end = atoms.find("</atom>", gt);
    double v[3];
    parse3(atoms.substr(gt + 1, end - gt - 1), v);
    int z = atomic_number(sym);
    if (!z) PPH_ERROR("unknown species '" + sym + "' in PROPhet.xml");
    sd.Z.push_back(z);
    cart.push_back(v[0]); cart.push_back(v[1]); cart.push_back(v[2]);
    from = end;
  }
  std::string tg = inner_last(b, "target");
  if (tg.empty()) PPH_ERROR("For system no target key");
  sd.target = atof(tg.c_str());
  finish_cartesian(sd, cart);
  return sd;
}

StructureData read_vasprun(const std::string& dir)
{
  const std::string filename = dir + "/vasprun.xml";
  std::ifstream f(filename.c_str());
  if (!f.is_open()) PPH_ERROR("Could not open " + filename);
  std::stringstream ss;
  ss << f.rdbuf();
  const std::string text = ss.str();
  StructureData sd;
  sd.train = "";
  auto last_varray = [&](const std::string& name) {
    size_t pos = std::string::npos, from = 0;
    for (;;) {
      size_t p = text.find("<varray name=\"" + name + "\"", from);
      if (p == std::string::npos) break;
      pos = p;
      from = p + 1;
    }
    if (pos == std::string::npos) PPH_ERROR("no varray '" + name + "' in " + filename);
    size_t end = text.find("</varray>", pos);
    return text.substr(pos, end - pos);
  };
  auto vectors = [&](const std::string& block) {
    std::vector<std::vector<double> > out;
    size_t from = 0;
    for (;;) {
      size_t p = block.find("<v>", from);
      if (p == std::string::npos) break;
      size_t e = block.find("</v>", p);
      std::istringstream s(block.substr(p + 3, e - p - 3));
      std::vector<double> v;
      double t;
      while (s >> t) v.push_back(t);
      out.push_back(v);
      from = e;
    }
    return out;
  };
  std::vector<std::vector<double> > basis = vectors(last_varray("basis"));
  if (basis.size() < 3) PPH_ERROR("basis not read from " + filename);
  for (int i = 0; i < 3; i++)
    for (int k = 0; k < 3; k++) sd.cell[3 * i + k] = basis[i][k];
  std::vector<std::vector<double> > pos = vectors(last_varray("positions"));
  size_t pa = text.find("<atoms>");
  int natom = pa == std::string::npos ? -1 : atoi(text.c_str() + pa + 7);
  if ((int)pos.size() != natom) PPH_ERROR("Atoms were not read correctly");
  // fractional coordinates taken as they are (CART = false: Normalize leaves them alone, Structure.cpp:1241-1242)
  for (size_t i = 0; i < pos.size(); i++)
    for (int k = 0; k < 3; k++) sd.pos.push_back(pos[i][k]);
  // atominfo: <rc><c>El</c><c>type</c></rc>; each entry repeated `type` times (VASP.cpp:139-143, kept as is)
  size_t ai = text.find("<array name=\"atoms\"");
  size_t ae = text.find("</array>", ai);
  std::string arr = text.substr(ai, ae - ai);
  std::vector<std::string> cs;
  size_t from = 0;
  for (;;) {
    size_t p = arr.find("<c>", from);
    if (p == std::string::npos) break;
    size_t e = arr.find("</c>", p);
    cs.push_back(arr.substr(p + 3, e - p - 3));
    from = e;
  }
  for (size_t i = 1; i < cs.size(); i += 2)
    for (int j = 0; j < atoi(cs[i].c_str()); j++) {
      int z = atomic_number(cs[i - 1]);
      if (!z) PPH_ERROR("unknown element '" + cs[i - 1] + "' in " + filename);
      sd.Z.push_back(z);
    }
  sd.Z.resize(pos.size(), sd.Z.empty() ? 0 : sd.Z.back());
  // energy: last <i name="e_wo_entrp"> (VASP.cpp:167-173)
  size_t pe = std::string::npos;
  from = 0;
  for (;;) {
    size_t p = text.find("<i name=\"e_wo_entrp\">", from);
    if (p == std::string::npos) break;
    pe = p;
    from = p + 1;
  }
  if (pe == std::string::npos) PPH_ERROR("Could not get energy from " + filename);
  sd.target = atof(text.c_str() + pe + 21);
  return sd;
}

// ---- Setup ----------------------------------------------------------------------------------------------------
static std::string to_lower(std::string s)
{
  std::transform(s.begin(), s.end(), s.begin(), ::tolower);
  return s;
}
static bool truthy(const std::string& v) { return v == "true" || v == "t" || v == "1"; }

std::string Setup::clean_line(std::string line)
{
  // Setup.cpp:691-717
  if (line.empty()) return line;
  for (size_t i = 0; i < line.size(); i++) {
    if (line[i] == '=') line.replace(i, 1, " ");
    if (line[i] == '#') line.resize(i);
  }
  if (!line.empty()) {
    const std::string ws = " \t\n\r";
    size_t a = line.find_first_not_of(ws);
    if (a != std::string::npos) line = line.substr(a, line.find_last_not_of(ws) - a + 1);
    else line.clear();
  }
  return line;
}

Setup::Setup(const std::string& input_file, Parallel* mpi_in, bool quiet) : mpi(mpi_in), my_is_potential(false)
{
  if (mpi->io_node() && !quiet) {
    std::cout << std::endl << "PROPhet (prophet_b200: B200 device path)" << std::endl << "run on:  ";
    time_t raw;
    time(&raw);
    std::cout << asctime(localtime(&raw)) << std::endl;
  }
  unsigned int seed = (unsigned int)time(NULL);
  if (mpi->io_node() && !quiet) std::cout << "random seed =  " << seed << std::endl;
  srand(seed);
  read_input(input_file);
  if (F.inputs.empty()) PPH_ERROR("No functional inputs given in input file");
  if (F.output.empty()) PPH_ERROR("No functional output given in input file");
  if (!my_is_potential) PPH_ERROR("only structure -> energy potentials are implemented (input must contain 'structure')");
  Ntotal_systems = (int)systems.size();
  split_systems();
  if (mpi->io_node() && !quiet) {
    // Setup::print_details (Setup.cpp:614-676)
    std::cout << std::endl << "Found " << Ntotal_systems << (Nsystems() == 1 ? " system" : " systems") << std::endl << std::endl;
    if (!F.FE.empty()) {                                                         // Setup.cpp:625-635
      std::cout << "############" << std::endl << "TRAINING TO FORMATION ENERGIES" << std::endl;
      std::cout << "Species Energies: " << std::endl;
      std::cout.precision(9);
      for (std::map<std::string, REAL>::const_iterator it = F.FE.begin(); it != F.FE.end(); ++it)
        std::cout << it->first << " : " << it->second << std::endl;
      std::cout.precision(6);
      std::cout << "############" << std::endl << std::endl;
    }
    std::cout << "Functional parameters" << std::endl;
    std::cout << F.inputs.size() << (F.inputs.size() == 1 ? " input" : " inputs") << " :  ";
    for (size_t i = 0; i < F.inputs.size(); i++) std::cout << F.inputs[i] << "   ";
    std::cout << std::endl;
    std::cout << F.hidden.size() << (F.hidden.size() == 1 ? " hidden layer" : " hidden layers") << " :  ";
    for (size_t i = 0; i < F.hidden.size(); i++) std::cout << F.hidden[i] << "   ";
    std::cout << std::endl << "output :  " << F.output << std::endl;
    std::cout << std::endl << "Running on " << mpi->Nprocs() << (mpi->Nprocs() == 1 ? " GPU" : " GPUs...") << std::endl << std::endl;
  }
}

void Setup::read_input(const std::string& filename_in)
{
  std::vector<std::ifstream*> files;
  std::ifstream* file = new std::ifstream(filename_in.c_str());
  if (!file->is_open()) PPH_ERROR("Could not open input file \"" + filename_in + "\"");
  files.push_back(file);
  std::string line, KEY, key, sv;
  int iv;
  REAL rv, Rcut_max = 0;
  std::getline(*file, line);
  while (true) {
    line = clean_line(line);
    if (files.front()->eof() && line.empty() && files.size() == 1) break;
    if (files.back()->eof() && files.size() > 1 && line.empty()) {
      files.back()->close();
      delete files.back();
      files.pop_back();
      std::getline(*files.back(), line);
      continue;
    }
    if (line.empty()) {
      std::getline(*files.back(), line);
      continue;
    }
    std::istringstream Line(line);
    Line >> KEY;
    key = to_lower(KEY);
    if (key == "data") {
      property_map.clear();
      while (Line >> KEY >> sv) {
        key = to_lower(KEY);
        if (key == "code") {
          property_map[key] = to_lower(sv);
          if (to_lower(sv) == "prophet") {
            // Setup.cpp:196-222: one system per <system> of PROPhet.xml in the CWD
            std::ifstream fx("PROPhet.xml");
            if (!fx.is_open()) PPH_ERROR("data code=prophet but PROPhet.xml is not in the working directory");
            std::string l;
            int nsys = -1;
            while (std::getline(fx, l)) {
              size_t p = l.find("<nsystem>");
              if (p != std::string::npos) { nsys = atoi(l.c_str() + p + 9); break; }
            }
            if (nsys < 0) PPH_ERROR("no <nsystem> line in PROPhet.xml");
            for (int i = 0; i < nsys; i++) {
              SystemSpec sp;
              std::ostringstream t;
              t << i;
              sp.files["base"] = t.str();
              for (std::map<std::string, std::string>::iterator it = property_map.begin(); it != property_map.end(); ++it) {
                if (it->first == "code") sp.files[it->first] = it->second;
                else sp.files[it->first] = sp.files["base"] + "/" + it->second;
              }
              systems.push_back(sp);
            }
          }
        } else {
          property_map[key] = sv;
        }
      }
    } else if (key == "input") {
      while (Line >> sv) {
        F.inputs.push_back(to_lower(sv));
        if (sv == "structure") my_is_potential = true;
      }
    } else if (key == "output") { Line >> sv; F.output = to_lower(sv); }
    else if (key == "hidden") { while (Line >> iv) if (iv != 0) F.hidden.push_back(iv); }
    else if (key == "transfer") { Line >> sv; F.transfer_function = to_lower(sv); }
    else if (key == "checkpoint_in") { Line >> sv; F.input_file = sv; }
    else if (key == "checkpoint_out") { Line >> sv; F.output_file = sv; }
    else if (key == "niterations") { Line >> iv; F.Niterations = iv; }
    else if (key == "ncheckpoint") { Line >> iv; F.Ncheckpoint = iv; }
    else if (key == "nprint") { Line >> iv; F.Nprint = iv; }
    else if (key == "nsave") { Line >> iv; F.Nbackup = iv; F.Tbackup = true; }
    else if (key == "ncycles") { Line >> iv; F.Ncycles = iv; }
    else if (key == "precondition_output") { Line >> sv; if (truthy(sv)) F.output_precondition = true; }
    else if (key == "precondition_input") { Line >> sv; F.input_precondition = truthy(sv); }
    else if (key == "print_gvector" || key == "print_gvectors") { Line >> sv; if (truthy(sv)) F.print_gvectors = true; }
    else if (key == "sgd") { Line >> sv; F.SGD = truthy(sv); }
    else if (key == "network_type") { Line >> sv; F.network_type = to_lower(sv); }
    else if (key == "algorithm") { Line >> sv; F.training_algorithm = to_lower(sv); }
    else if (key == "threshold") {
      Line >> rv;
      F.threshold = rv;
      if (F.line_min_epsilon == 0) {
        REAL eps100 = 100 * std::numeric_limits<REAL>::epsilon();
        F.line_min_epsilon = (0.1 * rv > eps100 ? 0.1 * rv : eps100);
      }
    }
    else if (key == "sd_momentum") { Line >> rv; F.sd_momentum = rv; }
    else if (key == "debug") { Line >> rv; F.debug = rv; }
    else if (key == "lambda_max") { Line >> rv; F.lambda_max = rv; }
    else if (key == "rcut") { Line >> rv; F.Rcut = rv; }
    else if (key == "nradial") { Line >> iv; F.Nradial = iv; }
    else if (key == "nangular") { Line >> iv; F.Nangular = iv; }
    else if (key == "regularization") { Line >> rv; F.regularization = rv; }
    else if (key == "line_min_epsilon") { Line >> rv; F.line_min_epsilon = rv; }
    else if (key == "dropout") { Line >> rv; F.dropout = rv; }
    else if (key == "t0") { Line >> rv; F.T0 = rv; }                          // Setup.cpp:479-503
    else if (key == "tf") { Line >> rv; F.Tf = rv; }
    else if (key == "nsteps_per_t") { Line >> iv; F.Nsteps_per_T = iv; }
    else if (key == "step_size") { Line >> rv; F.step_size = rv; }
    else if (key == "dt") { Line >> rv; F.dT = rv; }
    else if (key == "seed") { Line >> iv; srand(iv); }
    else if (key == "free_energy") { while (Line >> KEY >> rv) F.FE[KEY] = rv; }
    else if (key == "g1" || key == "g2" || key == "g3" || key == "g4") {
      std::vector<REAL> t;
      while (Line >> rv) t.push_back(rv);
      const size_t need = key == "g1" ? 2 : (key == "g2" ? 4 : 5);
      if (t.size() != need) {
        std::ostringstream m;
        m << "A " << KEY << " function requires " << need << " parameters";
        PPH_ERROR(m.str());
      }
      if (t[0] > Rcut_max) Rcut_max = t[0];
      (key == "g1" ? F.G1 : key == "g2" ? F.G2 : key == "g3" ? F.G3 : F.G4).push_back(t);
    }
    else if (key == "include") {
      Line >> sv;
      std::ifstream* nf = new std::ifstream(sv.c_str());
      if (!nf->is_open()) PPH_ERROR("Could not open nested input file '" + sv + "'");
      files.push_back(nf);
    }
    else if (key == "downsample" || key == "variance_bounds" || key == "conv_cd" || key == "conv_cd_stride" || key == "norm_cd" ||
             key == "t0" || key == "tf" || key == "dt" || key == "nsteps_per_t" || key == "step_size") {
      // density-functional / Monte-Carlo keys of workflows outside this path: accepted and ignored
    }
    else {
      // anything else must be a directory = one training example, optional trailing train|val|test (Setup.cpp:512-543)
      struct stat info;
      if (stat(KEY.c_str(), &info) != 0) PPH_ERROR("Unrecognized flag \"" + KEY + "\" in file \"" + filename_in + "\"");
      if (Rcut_max) F.Rcut = Rcut_max;
      sv.clear();
      std::string last;
      while (Line >> sv) last = sv;
      std::string train = (last == "train" || last == "val" || last == "test") ? last : "train";
      SystemSpec sp;
      sp.files["base"] = KEY;
      sp.files["train"] = train;
      for (std::map<std::string, std::string>::iterator it = property_map.begin(); it != property_map.end(); ++it) {
        if (it->first == "code") sp.files[it->first] = it->second;
        else sp.files[it->first] = sp.files["base"] + "/" + it->second;
      }
      systems.push_back(sp);
    }
    if (!std::getline(*files.back(), line)) line.clear();
  }
  for (size_t i = 0; i < files.size(); i++) { files[i]->close(); delete files[i]; }
}

void Setup::split_systems()
{
  // Setup.cpp:566-602: round robin i = rank, rank + Nprocs, ...
  std::vector<SystemSpec> mine;
  for (int i = mpi->rank(); i < (int)systems.size(); i += mpi->Nprocs()) {
    SystemSpec sp = systems[i];
    for (size_t j = 0; j < F.inputs.size(); j++)
      if (!sp.files.count(F.inputs[j])) sp.files[F.inputs[j]] = sp.files["base"];
    if (!sp.files.count(F.output)) sp.files[F.output] = sp.files["base"];
    sp.files["user"] = sp.files["base"];
    sp.files.erase("base");
    mine.push_back(sp);
  }
  systems.swap(mine);
}

REAL free_energy_sum(const std::map<std::string, REAL>& FE, const std::vector<int>& Z)
{
  REAL sum = 0.0;
  for (size_t a = 0; a < Z.size(); a++) {
    std::string sym = atomic_symbol(Z[a]);
    std::map<std::string, REAL>::const_iterator it = FE.find(sym);
    if (it == FE.end()) PPH_ERROR("No Free Energy supplied for " + sym);
    sum += it->second;
  }
  return sum;
}

StructureData Setup::load(int i) const
{
  // System::System (System.cpp:51-150)
  const SystemSpec& sp = systems.at(i);
  std::map<std::string, std::string>::const_iterator code = sp.files.find("code");
  if (code == sp.files.end()) PPH_ERROR("no 'data code=...' statement before the first system");
  StructureData sd;
  std::map<std::string, std::string>::const_iterator st = sp.files.find("structure");
  if (code->second == "prophet") {
    sd = read_prophet_xml(atoi(st->second.c_str()));
  } else if (code->second == "vasp") {
    sd = read_vasprun(st->second);
    if (F.output != "energy") PPH_ERROR("only output = energy is implemented for code=vasp");
  } else {
    PPH_ERROR("Interface to code '" + code->second + "' has not been implemented in this build (vasp and prophet are)");
  }
  std::map<std::string, std::string>::const_iterator tr = sp.files.find("train");
  std::string train = tr == sp.files.end() ? "" : tr->second;
  if (!sd.train.empty()) train = sd.train;      // System.cpp:116-119
  sd.train = train;
  // `free_energy` given: train on the total energy minus the per-atom reference energies (System.cpp:142-147 ->
  // Structure::train_Local, Structure.cpp:110-149)
  if (F.output == "energy" && !F.FE.empty()) sd.target -= free_energy_sum(F.FE, sd.Z);
  return sd;
}

}  // namespace pph
