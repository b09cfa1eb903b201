// main.cpp -- hypre_app: the mini-app driver over libb200hypre.  Usage: hypre_app INPUT_FILE (YAML).
// Mirrors /root/reference/src/main.cpp:31-229: device selection before HYPRE_Init, global HYPRE switches from
// solver_settings, `num_tests` repetitions of  setup_precon_and_solver -> load -> solve -> check_solution ->
// output_linear_system -> summarize_timers, "Total time" line, optional CSV of the timers (csv_profile_file).
// Ranks: one process per GPU, launched by torchrun or any launcher that sets RANK / WORLD_SIZE / LOCAL_RANK /
// MASTER_ADDR / MASTER_PORT (the MPI shim in include/mpi.h wires them over NCCL); a plain `hypre_app file.yaml` is one rank.
#include <chrono>
#include <cstdio>
#include <iostream>
#include <string>
#include <vector>

#include "HypreSystem.h"

int main(int argc, char *argv[]) {
  int iproc = 0, nproc = 1;
  int rc = 0;
  try {
    // picks the device from LOCAL_RANK before any allocation (main.cpp:41-45); no device => stop here, there is no CPU path
    if (MPI_Init(&argc, &argv) != MPI_SUCCESS) {
      fprintf(stderr, "hypre_app: cannot initialise the device / ranks: %s\n", b200hypre_last_error());
      return 1;
    }
    MPI_Comm_rank(MPI_COMM_WORLD, &iproc);
    MPI_Comm_size(MPI_COMM_WORLD, &nproc);
    {
      char name[256] = "";
      int device = 0, count = 0, major = 0, minor = 0;
      double free_gb = 0, total_gb = 0;
      b200hypre_device_info(name, (HYPRE_Int)sizeof(name), &device, &count, &major, &minor, &free_gb, &total_gb);
      if (iproc == 0)
        printf("\trank=%d : %s %s %d : %s (cc=%d.%d): device=%d of %d : free memory=%1.8g GB, total memory=%1.8g GB\n", iproc,
               __FUNCTION__, __FILE__, __LINE__, name, major, minor, device, count, free_gb, total_gb);
    }
    fflush(stdout);
    MPI_Barrier(MPI_COMM_WORLD);
    if (HYPRE_Init()) {
      fprintf(stderr, "hypre_app: HYPRE_Init failed: %s\n", b200hypre_last_error());
      return 1;
    }
    auto start = std::chrono::system_clock::now();
    if (argc != 2) {
      std::cout << "ERROR!! Incorrect arguments passed to program." << std::endl << "Usage: hypre_app INPUT_FILE" << std::endl << std::endl;
      return 1;
    }
    YAML::Node inpfile = YAML::LoadFile(argv[1]);
    YAML::Node node = inpfile["solver_settings"];

    HYPRE_SetMemoryLocation(HYPRE_MEMORY_DEVICE);
    HYPRE_SetExecutionPolicy(HYPRE_EXEC_DEVICE);
    struct { const char *key; const char *what; HYPRE_Int (*fn)(HYPRE_Int); } vendor[] = {
        {"spgemm_use_vendor", "SpGemm", HYPRE_SetSpGemmUseVendor},
        {"spmv_use_vendor", "SpMV", HYPRE_SetSpMVUseVendor},
        {"sptrans_use_vendor", "SpTrans", HYPRE_SetSpTransUseVendor},
    };
    for (auto &v : vendor) {   // accepted for input compatibility; the library always runs its own kernels
      int use = nalu::get_optional(node, v.key, 0);
      if (!iproc) std::cout << (use == 1 ? "Using VENDOR " : "NOT Using VENDOR ") << v.what << "." << std::endl;
      v.fn(use == 1);
    }

    // main.cpp:108-114 sizes hypre's Umpire device pool from this key; here it pre-grows the library's stream-ordered pool in one
    // piece (only when the input names it: the pool otherwise grows on demand and never shrinks).  fast_matrix_assemble /
    // fast_vector_assemble (HypreSystem.cpp:933,994) guard code the reference compiles out (#if 0): accepted, nothing to do.
    if (node["umpire_device_pool_mbs"]) {
      const long long mbs = node["umpire_device_pool_mbs"].as<long long>();
      if (!iproc) std::cout << "umpire_device_pool_mbs=" << mbs << std::endl;
      if (mbs > 0 && b200hypre_pool_reserve((unsigned long long)mbs * 1024ULL * 1024ULL))
        fprintf(stderr, "hypre_app: could not reserve %lld MiB of device pool: %s\n", mbs, b200hypre_last_error());
    }

    std::string csv_profile_file = nalu::get_optional<std::string>(node, "csv_profile_file", "");
    std::vector<std::string> names;
    std::vector<std::vector<double>> data;
    int num_tests = nalu::get_optional(node, "num_tests", 1);
    for (int i = 0; i < num_tests; ++i) {
      hypre_ResetDeviceRandGenerator(1234ULL, 0ULL);
      nalu::HypreSystem linsys(MPI_COMM_WORLD, inpfile);
      linsys.setup_precon_and_solver();
      linsys.checkMemory();
      linsys.load();
      linsys.checkMemory();
      linsys.solve();
      linsys.check_solution();
      linsys.output_linear_system();
      linsys.summarize_timers();
      if (!csv_profile_file.empty()) linsys.retrieve_timers(names, data);
      MPI_Barrier(MPI_COMM_WORLD);
      std::chrono::duration<double> elapsed = std::chrono::system_clock::now() - start;
      if (iproc == 0) std::cout << "Total time: " << elapsed.count() << " seconds" << std::endl;
      if (!linsys.all_close()) rc = 2;
      linsys.destroy_system();
    }
    if (!csv_profile_file.empty() && iproc == 0 && !names.empty()) {
      FILE *fid = fopen(csv_profile_file.c_str(), "wt");
      if (!fid) throw std::runtime_error("cannot write " + csv_profile_file);
      for (size_t i = 0; i < names.size(); ++i) fprintf(fid, "%s%s", names[i].c_str(), i + 1 < names.size() ? "," : "\n");
      for (size_t j = 0; j < data[0].size(); ++j)
        for (size_t i = 0; i < names.size(); ++i)
          fprintf(fid, "%1.15g%s", j < data[i].size() ? data[i][j] : 0.0, i + 1 < names.size() ? "," : "\n");
      fclose(fid);
    }
    HYPRE_Finalize();
    MPI_Finalize();
  } catch (const std::exception &e) {
    fprintf(stderr, "hypre_app (rank %d): %s\n", iproc, e.what());
    return 1;
  }
  return rc;
}
