// The step-35a program: same contract as the reference's main() (step-35a.cc:1712-1760) -- read
// "parameter-file.prm" from the current directory, build the solver, run it, return 1 with a banner on stderr if
// anything throws.  The reference fixes dimension and mesh at compile time (step-35a.cc:748,1726); here they are
// optional command-line arguments:
//
//     step-35a [dim [mesh-file [parameter-file]]]
#include "navier_stokes_projection.h"

#include <cstdlib>
#include <iostream>
#include <string>

namespace {

const char *const kRule = "----------------------------------------------------";

int report_failure(const char *what)
{
  std::cerr << "\n\n" << kRule << "\n";
  if (what)
    std::cerr << "Exception on processing: \n" << what << "\n";
  else
    std::cerr << "Unknown exception!\n";
  std::cerr << "Aborting!\n" << kRule << std::endl;
  return 1;
}

template <int dim>
void solve(const Step35::RunTimeParameters::Data_Storage &data, const std::string &mesh)
{
  Step35::NavierStokesProjection<dim> problem(data, mesh);
  problem.run(data.verbose, data.output_interval);
}

} // namespace

int main(int argc, char **argv)
{
  const int dim = argc > 1 ? std::atoi(argv[1]) : 3;
  const std::string mesh = argc > 2 ? argv[2] : "";
  const std::string prm = argc > 3 ? argv[3] : "parameter-file.prm";
  try
    {
      Step35::RunTimeParameters::Data_Storage data;
      data.read_data(prm.c_str());
      switch (dim)
        {
        case 2: solve<2>(data, mesh); break;
        case 3: solve<3>(data, mesh); break;
        default: throw std::runtime_error("the first argument (dimension) must be 2 or 3");
        }
    }
  catch (const std::exception &e)
    {
      return report_failure(e.what());
    }
  catch (...)
    {
      return report_failure(nullptr);
    }
  std::cout << kRule << "\nThe run finished without errors.\n" << std::endl;
  return 0;
}
