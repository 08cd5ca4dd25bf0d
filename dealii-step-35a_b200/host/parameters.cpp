#include "parameters.h"

#include <cmath>
#include <cstdlib>
#include <fstream>
#include <limits>
#include <sstream>
#include <stdexcept>
#include <vector>

namespace Step35 {
namespace RunTimeParameters {

namespace {
std::string trim(const std::string &s)
{
  const size_t b = s.find_first_not_of(" \t\r\n");
  if (b == std::string::npos)
    return "";
  const size_t e = s.find_last_not_of(" \t\r\n");
  return s.substr(b, e - b + 1);
}
const double kInf = std::numeric_limits<double>::infinity();
} // namespace

Data_Storage::Data_Storage()
{
  // names, defaults and patterns as declared at step-35a.cc:157-234
  declare("Method_Form", "rotational", Entry::Selection, 0, 0, "rotational|standard");
  declare("Physical data/initial_time", "0.", Entry::Double, 0., kInf);
  declare("Physical data/final_time", "1.", Entry::Double, 0., kInf);
  declare("Physical data/Reynolds", "1.", Entry::Double, 0., kInf);
  declare("Physical data/load_checkpoint_file_name", "", Entry::Anything);
  declare("Physical data/save_checkpoint_file_name", "", Entry::Anything);
  declare("Time step data/dt", "5e-4", Entry::Double, 0., kInf);
  declare("Space discretization/n_of_refines", "0", Entry::Integer, 0, 15);
  declare("Space discretization/pressure_fe_degree", "1", Entry::Integer, 1, 5);
  declare("Data solve velocity/max_iterations", "1000", Entry::Integer, 1, 1000);
  declare("Data solve velocity/eps", "1e-12", Entry::Double, 0., kInf);
  declare("Data solve velocity/Krylov_size", "30", Entry::Integer, 1, kInf);
  declare("Data solve velocity/off_diagonals", "60", Entry::Integer, 0, kInf);
  declare("Data solve velocity/diag_strength", "0.01", Entry::Double, 0., kInf);
  declare("Data solve velocity/update_prec", "15", Entry::Integer, 1, kInf);
  declare("verbose", "true", Entry::Bool);
  declare("output_interval", "1", Entry::Integer, 1, kInf);
  extract();
}

Data_Storage::~Data_Storage() {}

void Data_Storage::declare(const std::string &key, const std::string &def, Entry::Kind kind, double lo, double hi,
                           const std::string &choices)
{
  entries[key] = Entry{def, kind, lo, hi, choices};
}

void Data_Storage::set(const std::string &key, const std::string &value)
{
  auto it = entries.find(key);
  if (it == entries.end())
    throw std::runtime_error("parameter file: no entry with name <" + key + "> was declared");
  Entry &e = it->second;
  auto fail = [&]() {
    throw std::runtime_error("parameter file: the value <" + value + "> of entry <" + key + "> does not match its pattern");
  };
  switch (e.kind)
    {
    case Entry::Double:
    case Entry::Integer:
      {
        char *end = nullptr;
        const double v = std::strtod(value.c_str(), &end);
        if (end == value.c_str() || *end != '\0' || !(v >= e.lo && v <= e.hi))
          fail();
        if (e.kind == Entry::Integer && v != std::floor(v))
          fail();
        break;
      }
    case Entry::Bool:
      if (value != "true" && value != "false" && value != "yes" && value != "no")
        fail();
      break;
    case Entry::Selection:
      {
        bool found = false;
        std::stringstream ss(e.choices);
        std::string c;
        while (std::getline(ss, c, '|'))
          found = found || c == value;
        if (!found)
          fail();
        break;
      }
    case Entry::Anything: break;
    }
  e.value = value;
}

void Data_Storage::parse(std::istream &in)
{
  std::vector<std::string> path;
  std::string line;
  int lineno = 0;
  while (std::getline(in, line))
    {
      ++lineno;
      const size_t hash = line.find('#');
      if (hash != std::string::npos)
        line.erase(hash);
      line = trim(line);
      if (line.empty())
        continue;
      auto starts = [&](const char *w) {
        const std::string ws(w);
        return line.compare(0, ws.size(), ws) == 0 && (line.size() == ws.size() || line[ws.size()] == ' ' || line[ws.size()] == '\t');
      };
      if (starts("subsection") || starts("SUBSECTION"))
        path.push_back(trim(line.substr(10)));
      else if (starts("end") || starts("END"))
        {
          if (path.empty())
            throw std::runtime_error("parameter file line " + std::to_string(lineno) + ": 'end' without subsection");
          path.pop_back();
        }
      else if (starts("set") || starts("SET"))
        {
          const size_t eq = line.find('=');
          if (eq == std::string::npos)
            throw std::runtime_error("parameter file line " + std::to_string(lineno) + ": expected 'set name = value'");
          const std::string name = trim(line.substr(3, eq - 3));
          const std::string value = trim(line.substr(eq + 1));
          std::string key;
          for (const auto &p : path)
            key += p + "/";
          set(key + name, value);
        }
      else
        throw std::runtime_error("parameter file line " + std::to_string(lineno) + ": cannot parse <" + line + ">");
    }
  if (!path.empty())
    throw std::runtime_error("parameter file: subsection <" + path.back() + "> is not closed");
}

void Data_Storage::extract()
{
  auto str = [&](const char *k) { return entries.at(k).value; };
  auto dbl = [&](const char *k) { return std::strtod(entries.at(k).value.c_str(), nullptr); };
  auto uns = [&](const char *k) { return (unsigned int)std::strtod(entries.at(k).value.c_str(), nullptr); };
  form = str("Method_Form") == "rotational" ? METHOD_ROTATIONAL : METHOD_STANDARD;
  initial_time = dbl("Physical data/initial_time");
  final_time = dbl("Physical data/final_time");
  Reynolds = dbl("Physical data/Reynolds");
  load_checkpoint_file_name = str("Physical data/load_checkpoint_file_name");
  save_checkpoint_file_name = str("Physical data/save_checkpoint_file_name");
  dt = dbl("Time step data/dt");
  n_global_refines = uns("Space discretization/n_of_refines");
  pressure_degree = uns("Space discretization/pressure_fe_degree");
  vel_max_iterations = uns("Data solve velocity/max_iterations");
  vel_eps = dbl("Data solve velocity/eps");
  vel_Krylov_size = uns("Data solve velocity/Krylov_size");
  vel_off_diagonals = uns("Data solve velocity/off_diagonals");
  vel_diag_strength = dbl("Data solve velocity/diag_strength");
  vel_update_prec = uns("Data solve velocity/update_prec");
  const std::string v = str("verbose");
  verbose = (v == "true" || v == "yes");
  output_interval = uns("output_interval");
}

void Data_Storage::read_data(const char *filename)
{
  std::ifstream file(filename);
  if (!file)
    throw std::runtime_error(std::string("cannot open parameter file <") + filename + ">");
  parse(file);
  extract();
}

void Data_Storage::read_data_from_string(const std::string &text)
{
  std::istringstream in(text);
  parse(in);
  extract();
}

} // namespace RunTimeParameters
} // namespace Step35
