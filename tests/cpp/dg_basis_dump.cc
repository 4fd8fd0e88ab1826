// Prints the 1-D tables of the product's DG bases (multigrid_b200/csrc/basis.h, host code shared with the CUDA
// library) so that the CPU test suite can compare them with the oracle's independent numpy construction.
// usage: dg_basis_dump <degree> <basis>      one line per table: name followed by the values
#include <cstdio>
#include <cstdlib>

#include "basis.h"

static void
dump(const char *name, const std::vector<double> &v)
{
  std::printf("%s", name);
  for (const double x : v)
    std::printf(" %.17e", x);
  std::printf("\n");
}

int
main(int argc, char **argv)
{
  if (argc != 3)
    return 2;
  const mgb::DGBasis b = mgb::make_dg_basis(std::atoi(argv[1]), std::atoi(argv[2]));
  dump("S", b.S);
  dump("G", b.G);
  dump("D", b.D);
  dump("f0", b.f0);
  dump("f1", b.f1);
  dump("e0", b.e0);
  dump("e1", b.e1);
  dump("B", b.B);
  dump("Binv", b.Binv);
  dump("xq", b.xq);
  dump("wq", b.wq);
  return 0;
}
