// tests/shim_host_check.cpp -- host-only utilities of include/shim/FluidSolverGPU.hpp (Monitor, should_save,
// get_output_directory, init_output_directory) exercised without a GPU; tests/test_output_cpu.py compares the
// monitor log and the should_save sequence with what the unmodified reference produces for the same input.
//   shim_host_check <monitor.log> <dir to create>   reads rows "f0 f1 f2 f3 i0 i1" from stdin
#include <cstdio>
#include <iostream>

#include "shim/FluidSolverGPU.hpp"

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  double f[4] = {0, 0, 0, 0};
  Index iv[2] = {0, 0};
  {
    Monitor<double> monitor(argv[1]);
    // the column order of oracle/ref_harness.cpp: fsref_monitor_log
    monitor.add_variable(&f[0], "time");
    monitor.add_variable(&f[1], "dt");
    monitor.add_variable(&f[2], "max(U)");
    monitor.add_variable(&iv[0], "iter(p)");
    monitor.add_variable(&iv[1], "a rather long column name");
    monitor.add_variable(&f[3], "momentum (x)");
    long long a, b;
    while (std::cin >> f[0] >> f[1] >> f[2] >> f[3] >> a >> b) {
      iv[0] = a;
      iv[1] = b;
      monitor.write();
    }
  }
  // should_save over a fixed sequence of steps (dt_write = 1, t_end = 3)
  double t = 0.0;
  const double dts[] = {0.3, 0.3, 0.3, 0.3, 0.35, 0.45, 0.5, 0.25, 0.25};
  std::printf("save");
  for (double dt : dts) {
    t += dt;
    std::printf(" %d", should_save(t, dt, 1.0, 3.0) ? 1 : 0);
  }
  std::printf(" %d\n", should_save(3.0, 0.25, 1.0, 3.0) ? 1 : 0);
  std::printf("dir %s\n", get_output_directory("test/output", "/base").c_str());
  const std::string d = std::string(argv[2]) + "/a/b";
  if (!init_output_directory(d)) return 3;
  if (std::FILE* fp = std::fopen((d + "/old.txt").c_str(), "w")) std::fclose(fp);
  if (!init_output_directory(d)) return 3;
  std::printf("emptied %d\n", std::filesystem::is_empty(d) ? 1 : 0);
  return 0;
}
