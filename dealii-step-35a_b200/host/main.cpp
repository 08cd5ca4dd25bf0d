// The step-35a program (reference main(), step-35a.cc:1712-1760): reads
// "parameter-file.prm" from the current directory, builds the solver and runs it.
// The reference fixes dimension and mesh at compile time (step-35a.cc:748,1726);
// here they can be given on the command line:  step-35a [dim [mesh-file [prm-file]]]
#include "navier_stokes_projection.h"

#include <cstdlib>
#include <iostream>

int main(int argc, char **argv)
{
  try
    {
      using namespace Step35;
      const int dim = argc > 1 ? std::atoi(argv[1]) : 3;
      const std::string mesh = argc > 2 ? argv[2] : "";
      const std::string prm = argc > 3 ? argv[3] : "parameter-file.prm";
      RunTimeParameters::Data_Storage data;
      data.read_data(prm.c_str());
      if (dim == 2)
        {
          NavierStokesProjection<2> test(data, mesh);
          test.run(data.verbose, data.output_interval);
        }
      else if (dim == 3)
        {
          NavierStokesProjection<3> test(data, mesh);
          test.run(data.verbose, data.output_interval);
        }
      else
        throw std::runtime_error("dim must be 2 or 3");
    }
  catch (std::exception &exc)
    {
      std::cerr << std::endl
                << std::endl
                << "----------------------------------------------------" << std::endl;
      std::cerr << "Exception on processing: " << std::endl
                << exc.what() << std::endl
                << "Aborting!" << std::endl
                << "----------------------------------------------------" << std::endl;
      return 1;
    }
  catch (...)
    {
      std::cerr << std::endl
                << std::endl
                << "----------------------------------------------------" << std::endl;
      std::cerr << "Unknown exception!" << std::endl
                << "Aborting!" << std::endl
                << "----------------------------------------------------" << std::endl;
      return 1;
    }
  std::cout << "----------------------------------------------------" << std::endl
            << "Apparently everything went fine!" << std::endl
            << "Don't forget to brush your teeth :-)" << std::endl
            << std::endl;
  return 0;
}
