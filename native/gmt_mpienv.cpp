// gmt_mpienv -- native counterpart of mpienv (mpienv.f90:1-35): every rank prints the integer it
// reads from the environment variable MEMORY_PER_CORE (0 when unset or not a number), the
// reference's aid for launchers that swallow environment variables.  One process plays all ranks
// (--ranks R, default: the number of GPUs, 1 without a GPU).
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "gmt_b200.h"

int main(int argc, char** argv)
{
  int ranks = 0;
  for (int i = 1; i < argc; ++i)
    if (!std::strcmp(argv[i], "--ranks") && i + 1 < argc) ranks = std::atoi(argv[++i]);
  if (ranks <= 0) {
    int n_dev = 0;
    if (gmt_device_count(&n_dev) != 0 || n_dev <= 0) n_dev = 1;
    ranks = n_dev;
  }
  const char* raw = std::getenv("MEMORY_PER_CORE");
  const int value = raw ? std::atoi(raw) : 0;
  for (int r = 0; r < ranks; ++r) std::printf(" rank %11d  MEMORY_PER_CORE=%11d\n", r, value);
  return 0;
}
