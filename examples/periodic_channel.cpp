// examples/periodic_channel.cpp -- the reference's test/PeriodicChannel.cpp (constants :18-52, time loop
// :160-262, pass criteria :264-337) written against include/shim/FluidSolverGPU.hpp: same calls, same order,
// the same tests at the end; every operation runs on the B200 through libfsgpu.so.  The driver-inlined host
// loops of the reference become update_velocity / force_total_flow / correct_velocity (see the shim header).
//
//   ./examples/periodic_channel [T_END = 60] [vtk output directory]     exit code 0 = all of the reference's tests pass
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "shim/FluidSolverGPU.hpp"

using Float = double;
constexpr Index NX = 5 * 43, NY = 43, NGHOST = 1;
constexpr Float X_MIN = 0.0, X_MAX = 5.0, Y_MIN = 0.0, Y_MAX = 1.0;
constexpr Float DT_MAX = 1e-1, CFL_MAX = 0.9, DT_WRITE = 1.0;
constexpr Float U_INIT = 1.0, VISC = 1e-3, RHO = 0.5;
constexpr Float TOTAL_FLOW      = (Y_MAX - Y_MIN) * U_INIT * RHO;
constexpr int PRESSURE_MAX_ITER = 50;
constexpr Float PRESSURE_TOL    = 1e-6;
constexpr Index NUM_SUBITER     = 2;

// src/Quadrature.hpp:76-90
static Float trapezoidal_rule(const std::vector<Float>& f, const std::vector<Float>& x) {
  Float integral = 0.0;
  for (size_t i = 0; i + 1 < f.size(); ++i) integral += (x[i + 1] - x[i]) * 0.5 * (f[i + 1] + f[i]);
  return integral;
}

int main(int argc, char** argv) {
  const Float T_END         = argc > 1 ? std::atof(argv[1]) : 60.0;
  const std::string vtk_dir = argc > 2 ? argv[2] : "";
  const FlowBConds<Float> bconds{
      .left   = Periodic{},
      .right  = Periodic{},
      .bottom = Dirichlet<Float>{.U = 0.0, .V = 0.0},
      .top    = Dirichlet<Float>{.U = 0.0, .V = 0.0},
  };
  FS<Float, NX, NY, NGHOST> fs{.visc_gas = VISC, .visc_liquid = VISC, .rho_gas = RHO, .rho_liquid = RHO};
  init_grid(X_MIN, X_MAX, NX, Y_MIN, Y_MAX, NY, fs);
  calc_rho(fs);
  calc_visc(fs);

  Field2D<Float, NX, NY, NGHOST> Ui{}, Vi{}, div{}, delta_p{};
  Field2D<Float, NX + 1, NY, NGHOST> drhoUdt{};
  Field2D<Float, NX, NY + 1, NGHOST> drhoVdt{};

  LinearSolver_StructHypre<Float, NX, NY, NGHOST> ps(PRESSURE_TOL, PRESSURE_MAX_ITER);
  ps.set_pressure_operator(fs);

  for_each_i<Exec::Parallel>(fs.curr.U, [&](Index i, Index j) { fs.curr.U(i, j) = U_INIT; });
  for_each_i<Exec::Parallel>(fs.curr.V, [&](Index i, Index j) { fs.curr.V(i, j) = 0.0; });
  fs.curr.U.upload();
  fs.curr.V.upload();
  apply_velocity_bconds(fs, bconds);

  DataWriter<Float, NX, NY, NGHOST> data_writer(vtk_dir, &fs.x, &fs.y);
  data_writer.add_scalar("pressure", &fs.p);
  data_writer.add_scalar("divergence", &div);
  data_writer.add_vector("velocity", &Ui, &Vi);

  Float t = 0.0, dt = DT_MAX, inflow = 0, outflow = 0, p_res = 0, next_write = 0.0;
  Index steps = 0, p_iter_total = 0;
  bool any_test_failed = false;
  while (t < T_END) {
    dt = adjust_dt(fs, CFL_MAX, DT_MAX);
    dt = std::min(dt, T_END - t);
    save_old_state(fs.curr, fs.old);
    for (Index sub_iter = 0; sub_iter < NUM_SUBITER; ++sub_iter) {
      calc_mid_time(fs.curr.U, fs.old.U);
      calc_mid_time(fs.curr.V, fs.old.V);
      calc_dmomdt(fs, drhoUdt, drhoVdt);
      update_velocity(drhoUdt, drhoVdt, dt, fs);
      apply_velocity_bconds(fs, bconds);
      force_total_flow(fs, TOTAL_FLOW, inflow, outflow);
      calc_divergence(fs.curr.U, fs.curr.V, fs.dx, fs.dy, div);
      Index local_p_iter = 0;
      ps.set_pressure_rhs(fs, div, dt);
      ps.solve(delta_p, &p_res, &local_p_iter);
      p_iter_total += local_p_iter;
      shift_pressure_to_zero(fs.dx, fs.dy, delta_p);
      correct_velocity(fs, delta_p, dt);
    }
    t += dt;
    ++steps;
    {  // test/PeriodicChannel.cpp:236-258: the two ghost columns still carry TOTAL_FLOW after the projection
      Float in2 = 0, out2 = 0;
      fs.curr.U.download();
      for (Index j = -NGHOST; j < NY + NGHOST; ++j) {
        in2  += RHO * fs.curr.U(-NGHOST, j) * fs.dy;
        out2 += RHO * fs.curr.U(NX + NGHOST, j) * fs.dy;
      }
      if (std::abs(in2 - TOTAL_FLOW) > 1e-8 || std::abs(out2 - in2) > 1e-8) {
        std::printf("t=%.6e: inflow %.6e outflow %.6e TOTAL_FLOW %.6e\n", t, in2, out2, TOTAL_FLOW);
        any_test_failed = true;
      }
    }
    if (!vtk_dir.empty() && t >= next_write) {
      interpolate_U(fs.curr.U, Ui);
      interpolate_V(fs.curr.V, Vi);
      calc_divergence(fs.curr.U, fs.curr.V, fs.dx, fs.dy, div);
      if (!data_writer.write(t)) return 1;
      next_write += DT_WRITE;
    }
  }
  calc_divergence(fs.curr.U, fs.curr.V, fs.dx, fs.dy, div);
  Float mass = 0, mom_x = 0, mom_y = 0;
  calc_conserved_quantities(fs, mass, mom_x, mom_y);
  std::printf("steps %lld t %.6e U_max %.12e V_max %.6e div_max %.3e p_iter/solve %.2f mass %.12e mom_x %.12e\n",
              (long long)steps, t, abs_max(fs.curr.U), abs_max(fs.curr.V), abs_max(div),
              (double)p_iter_total / (double)(steps * NUM_SUBITER), mass, mom_x);

  // = Perform tests (test/PeriodicChannel.cpp:264-337) ==============================================
  fs.to_host();
  {
    constexpr Float TOL = 1e-4;
    for (Index i = 0; i < NX; ++i)
      for (Index j = 0; j < NY; ++j)
        if (std::abs(fs.p(i, j) - fs.p(i, 0)) > TOL) {
          std::printf("Non constant pressure along y-axis for x=%g\n", fs.xm(i));
          any_test_failed = true;
          break;
        }
    const Float ref_dpdx = (fs.p(NX / 2 + 1, NY / 2) - fs.p(NX / 2, NY / 2)) / fs.dx;
    for (Index i = 1; i < NX; ++i) {
      const Float dpdx = (fs.p(i, NY / 2) - fs.p(i - 1, NY / 2)) / fs.dx;
      if (std::abs(ref_dpdx - dpdx) > TOL) {
        std::printf("Non constant dpdx: %.6e at x=%g, reference %.6e\n", dpdx, fs.x(i), ref_dpdx);
        any_test_failed = true;
      }
    }
  }
  {
    constexpr Float TOL = 1e-3;
    std::vector<Float> diff(NY + 2 * NGHOST), ym(NY + 2 * NGHOST);
    for (Index i = 0; i < NX + 1; ++i) {
      for (Index j = -NGHOST; j < NY + NGHOST; ++j) {
        const Float dpdx   = (fs.p(i, j) - fs.p(i - 1, j)) / fs.dx;
        const Float y      = fs.ym(j);
        ym[j + NGHOST]     = y;
        diff[j + NGHOST]   = std::abs(fs.curr.U(i, j) - dpdx / (2 * VISC) * (y * y - y));
      }
      const Float L1_error = trapezoidal_rule(diff, ym);
      if (L1_error > TOL) {
        std::printf("U profile at x=%g: L1 error %.6e\n", fs.x(i), L1_error);
        any_test_failed = true;
      }
    }
  }
  {
    constexpr Float TOL = 1e-7;
    for (Index i = 0; i < NX; ++i)
      for (Index j = 0; j < NY + 1; ++j)
        if (std::abs(fs.curr.V(i, j)) > TOL) {
          std::printf("V(%g, %g) = %.6e is not zero\n", fs.xm(i), fs.y(j), fs.curr.V(i, j));
          any_test_failed = true;
        }
  }
  std::printf(any_test_failed ? "FAILED\n" : "all tests of test/PeriodicChannel.cpp passed\n");
  return any_test_failed ? 1 : 0;
}
