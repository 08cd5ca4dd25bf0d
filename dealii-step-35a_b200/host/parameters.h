// RunTimeParameters::Data_Storage of the reference (step-35a.cc:121-293): same
// field names, same entry names, defaults and admissible ranges, same file
// syntax (`set key = value`, `subsection name` ... `end`, `#` comments).
// deal.II's ParameterHandler is replaced by a small parser.
#pragma once
#include <map>
#include <string>

namespace Step35 {
namespace RunTimeParameters {

enum MethodFormulation
{
  METHOD_STANDARD,
  METHOD_ROTATIONAL
};

class Data_Storage
{
public:
  Data_Storage();
  ~Data_Storage();
  // throws std::runtime_error if the file cannot be opened, names an undeclared
  // entry, or holds a value outside the declared pattern
  void read_data(const char *filename);
  // same, from a string holding the file's text
  void read_data_from_string(const std::string &text);

  MethodFormulation form;
  double initial_time, final_time, Reynolds;
  std::string load_checkpoint_file_name;
  std::string save_checkpoint_file_name;
  double dt;
  unsigned int n_global_refines, pressure_degree;
  unsigned int vel_max_iterations, vel_Krylov_size, vel_off_diagonals, vel_update_prec;
  double vel_eps, vel_diag_strength;
  bool verbose;
  unsigned int output_interval;

protected:
  struct Entry
  {
    std::string value;
    enum Kind
    {
      Double,
      Integer,
      Bool,
      Selection,
      Anything
    } kind;
    double lo, hi; // admissible range for Double / Integer
    std::string choices;
  };
  // key = "subsection/name" (or "name" at top level)
  std::map<std::string, Entry> entries;
  void declare(const std::string &key, const std::string &def, Entry::Kind kind, double lo = 0, double hi = 0,
               const std::string &choices = "");
  void set(const std::string &key, const std::string &value);
  void parse(std::istream &in);
  void extract();
};

} // namespace RunTimeParameters
} // namespace Step35
