// bench_dolfinx -- the reference's command line (src/main.cpp:136-320) on the B200-native library.
//
// Same options, defaults, console lines and JSON keys.  Differences: there is no MPI in this
// environment, so ranks are host threads of one process, one per GPU (--ngpus N, default 1), joined
// by NCCL; --platform accepts only "gpu" (there is no CPU fallback); Boost is replaced by a small
// parser that, like the reference's allow_unregistered(), ignores options it does not know.
#include <barrier>
#include <cstdio>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>
#include <thread>
#include <vector>

#include "laplacian_solver.hpp"

using namespace benchdolfinx;

namespace
{
struct Options
{
  std::string platform = "gpu";
  std::size_t float_size = 64, ndofs = 1000, ndofs_global = 0, qmode = 1, nreps = 1000, degree = 3;
  bool cg = false, mat_comp = false, use_gauss = false, help = false;
  bool ndofs_set = false, ndofs_global_set = false;
  double geom_perturb_fact = 0.0;
  std::string json;
  int ngpus = 1;
};

void usage()
{
  std::cout << "DOLFINx benchmark\n-----------------\n";
  std::cout << "\n  Finite Element Operator Action Benchmark which computes\n";
  std::cout << "  the Laplacian operator on a cube mesh of hexahedral elements.\n\n";
  std::cout << "Options:\n"
               "  -h [ --help ]                   Print usage message\n"
               "  --platform arg (=gpu)           Compute platform (gpu only: no CPU fallback)\n"
               "  --float arg (=64)               Float size (bits). 32 or 64.\n"
               "  --ndofs arg (=1000)             Number of degrees-of-freedom per rank (GPU)\n"
               "  --ndofs_global arg (=0)         Number of global degrees-of-freedom\n"
               "  --qmode arg (=1)                Quadrature mode (0 or 1): qmode=0 has P+1 points in\n"
               "                                  each direction, qmode=1 has P+2 points in each direction.\n"
               "  --cg                            Do CG iterations, rather than simple operator action\n"
               "  --nreps arg (=1000)             Number of repetitions\n"
               "  --degree arg (=3)               Polynomial degree \"P\" (1-7)\n"
               "  --mat_comp                      Compare result to matrix operator (slow with large ndofs)\n"
               "  --geom_perturb_fact arg (=0)    Randomly perturb the geometry (useful to check correctness)\n"
               "  --use_gauss                     Use Gauss quadrature rather than GLL quadrature\n"
               "  --json arg                      Filename for JSON output\n"
               "  --ngpus arg (=1)                Number of ranks = GPUs of this node (replaces mpirun -n)\n";
}

Options parse(int argc, char** argv)
{
  Options o;
  auto value = [&](int& i, const std::string& arg, const std::string& name, std::string& out) {
    if (arg == name && i + 1 < argc)
    {
      out = argv[++i];
      return true;
    }
    if (arg.rfind(name + "=", 0) == 0)
    {
      out = arg.substr(name.size() + 1);
      return true;
    }
    return false;
  };
  for (int i = 1; i < argc; ++i)
  {
    std::string a = argv[i], v;
    if (a == "-h" || a == "--help")
      o.help = true;
    else if (a == "--cg")
      o.cg = true;
    else if (a == "--mat_comp")
      o.mat_comp = true;
    else if (a == "--use_gauss")
      o.use_gauss = true;
    else if (value(i, a, "--platform", v))
      o.platform = v;
    else if (value(i, a, "--float", v))
      o.float_size = std::stoul(v);
    else if (value(i, a, "--ndofs_global", v))
    {
      o.ndofs_global = std::stoull(v);
      o.ndofs_global_set = true;
    }
    else if (value(i, a, "--ndofs", v))
    {
      o.ndofs = std::stoull(v);
      o.ndofs_set = true;
    }
    else if (value(i, a, "--qmode", v))
      o.qmode = std::stoul(v);
    else if (value(i, a, "--nreps", v))
      o.nreps = std::stoul(v);
    else if (value(i, a, "--degree", v))
      o.degree = std::stoul(v);
    else if (value(i, a, "--geom_perturb_fact", v))
      o.geom_perturb_fact = std::stod(v);
    else if (value(i, a, "--json", v))
      o.json = v;
    else if (value(i, a, "--ngpus", v))
      o.ngpus = std::stoi(v);
    // unknown options are ignored (allow_unregistered, src/main.cpp:187-190)
  }
  if (o.ndofs_set && o.ndofs_global_set)
    throw std::logic_error("Conflicting options 'ndofs' and 'ndofs_global'"); // src/main.cpp:192-196
  return o;
}

std::string json_number(double v)
{
  char buf[64];
  std::snprintf(buf, sizeof(buf), "%.16E", v);
  return buf;
}

template <typename T>
BenchmarkResults run_benchmark(const Options& o, std::array<std::int64_t, 3> nx, int rank, int size, bd_comm* comm,
                               std::barrier<>& sync, std::int64_t& ncells_global, std::int64_t& ndofs_global)
{
  BoxMesh<T> mesh(nx, static_cast<int>(o.degree), rank, size, o.geom_perturb_fact, comm);
  ncells_global = mesh.info.ncells_global;
  ndofs_global = mesh.info.ndofs_global;
  return laplace_action_gpu<T>(mesh, static_cast<int>(o.degree), static_cast<int>(o.qmode), T(2.0),
                               static_cast<int>(o.nreps), o.use_gauss, o.mat_comp, o.cg, rank,
                               [&] { sync.arrive_and_wait(); });
}
} // namespace

int main(int argc, char* argv[])
{
  Options o;
  try
  {
    o = parse(argc, argv);
  }
  catch (const std::exception& e)
  {
    std::cerr << e.what() << std::endl;
    return 1;
  }
  if (o.help)
  {
    usage();
    return 0;
  }
  try
  {
    if (o.float_size != 32 && o.float_size != 64)
      throw std::runtime_error("Invalid float size. Must be 32 or 64.");
    if (o.qmode > 1)
      throw std::runtime_error("Invalid qmode.");
    if (o.platform != "gpu")
      throw std::runtime_error("Invalid platform: " + o.platform + " (this build has no CPU fallback)");
    const int size = o.ngpus;
    int ndev = 0;
    check_cuda(cudaGetDeviceCount(&ndev));
    if (size < 1 || size > ndev)
      throw std::runtime_error("--ngpus must be between 1 and the number of visible GPUs");
    std::size_t ndofs = o.ndofs, ndofs_global = o.ndofs_global;
    if (!o.ndofs_global_set)
      ndofs_global = ndofs * size; // src/main.cpp:237-240
    else
      ndofs = ndofs_global / size;

    char banner[1024];
    check(bd_device_information(banner, sizeof(banner)));
    std::cout << banner;
    std::cout << "-----------------------------------\n";
    std::cout << "Platform: " << o.platform << "\n";
    std::cout << "Polynomial degree : " << o.degree << "\n";
    std::cout << "Number of ranks : " << size << std::endl;
    std::cout << "Requested number of local DoFs : " << ndofs << std::endl;
    std::cout << "Number of repetitions : " << o.nreps << std::endl;
    std::cout << "Scalar Type: " << o.float_size << std::endl;
    std::cout << "Use Gauss-Jacobi: " << o.use_gauss << std::endl;
    std::cout << "Compare to matrix: " << o.mat_comp << std::endl;
    std::cout << "-----------------------------------" << std::endl;

    const auto nx = compute_mesh_size(static_cast<std::int64_t>(ndofs_global), static_cast<int>(o.degree));

    char id[BD_UNIQUE_ID_BYTES] = {0};
    if (size > 1)
      check(bd_comm_unique_id(id));
    std::barrier<> sync(size);
    std::vector<BenchmarkResults> results(size);
    std::vector<std::string> errors(size);
    std::int64_t ncells_global = 0, ndofs_global_out = 0;
    auto worker = [&](int rank) {
      try
      {
        check_cuda(cudaSetDevice(rank));
        bd_comm* comm = nullptr;
        check(bd_comm_create(size > 1 ? id : nullptr, rank, size, &comm));
        std::int64_t nc = 0, ndg = 0;
        results[rank] = o.float_size == 32 ? run_benchmark<float>(o, nx, rank, size, comm, sync, nc, ndg)
                                           : run_benchmark<double>(o, nx, rank, size, comm, sync, nc, ndg);
        if (rank == 0)
        {
          ncells_global = nc;
          ndofs_global_out = ndg;
        }
        bd_comm_destroy(comm);
      }
      catch (const std::exception& e)
      {
        errors[rank] = e.what();
        if (size > 1)
        {
          // The other ranks are (or will be) inside NCCL collectives that this rank will never join: leaving the
          // barrier would let them hang.  Same outcome as an MPI rank of the reference dying in err_check
          // (src/util.hpp:36-45: printf + exit(1)): the whole job ends.
          std::fprintf(stderr, "rank %d: %s\n", rank, e.what());
          std::fflush(stderr);
          std::_Exit(1);
        }
        sync.arrive_and_drop();
      }
    };
    std::vector<std::thread> threads;
    for (int r = 1; r < size; ++r)
      threads.emplace_back(worker, r);
    worker(0);
    for (auto& t : threads)
      t.join();
    for (int r = 0; r < size; ++r)
      if (!errors[r].empty())
        throw std::runtime_error("rank " + std::to_string(r) + ": " + errors[r]);

    const BenchmarkResults& res = results[0];
    if (!o.json.empty())
    {
      // same keys as src/main.cpp:262-270,122-130
      std::ostringstream js;
      js << "{\"input\":{\"p\":" << o.degree << ",\"mpi_size\":" << size << ",\"ndofs_local_requested\":" << ndofs
         << ",\"nreps\":" << o.nreps << ",\"scalar_size\":" << o.float_size
         << ",\"use_gauss\":" << (o.use_gauss ? "true" : "false") << ",\"mat_comp\":" << (o.mat_comp ? "true" : "false")
         << ",\"qmode\":" << o.qmode << ",\"cg\":" << (o.cg ? "true" : "false") << "},\"output\":{\"ncells_global\":"
         << ncells_global << ",\"ndofs_global\":" << ndofs_global_out
         << ",\"mat_free_time\":" << json_number(res.mat_free_time) << ",\"u_norm\":" << json_number(res.unorm)
         << ",\"y_norm\":" << json_number(res.ynorm) << ",\"z_norm\":" << json_number(res.znorm)
         << ",\"gdof_per_second\":" << json_number(ndofs_global_out * o.nreps / (1e9 * res.mat_free_time)) << "}}";
      std::filesystem::path filename(o.json);
      std::cout << "*** Writing output to:       " << filename << std::endl;
      std::cout << "*** Writing output to (abs): " << std::filesystem::absolute(filename) << std::endl;
      std::ofstream strm(filename, std::ofstream::out);
      strm << js.str() << "\n";
    }
    else
      std::cout << "*** Empty file: " << o.json << std::endl;
  }
  catch (const std::exception& e)
  {
    std::cerr << "bench_dolfinx: " << e.what() << std::endl;
    return 1;
  }
  return 0;
}
