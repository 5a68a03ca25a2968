// cudaMpiKNN_prePartitionedData — one spatially coherent float3 file per rank (listed in a text
// file), one output file per rank named <prefix>_%06d.float. Same command line and formats as the
// reference program of the same name (/root/reference/prePartitionedDataVariant.cu:176-388):
//   knnrun -n <numFiles> ./cudaMpiKNN_prePartitionedData fileNames.txt -o prefix -k 100 [-r R] [-g G]
#include "knn_apps.hpp"

using namespace knnapp;

int main(int ac, char** av) {
  try {
    const Options o = parse_args(ac, av, "input file name (should be a text file with list of input files)",
                                 "output file(s) prefix");
    const std::vector<std::string> names = read_file_list(o.input);
    Bootstrap bs;
    if ((size_t)bs.size() != names.size())
      throw std::runtime_error("number of input files does not match number of ranks");
    std::vector<float> pts = read_points(names[bs.rank()], 0, 1, nullptr, nullptr);
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (size_t i = 0; i + 2 < pts.size(); i += 3)
      for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], pts[i + a]); hi[a] = fmaxf(hi[a], pts[i + a]); }
    std::cout << "#" << bs.rank() << "/" << bs.size() << ": got " << pts.size() / 3 << " points to work on, bounds is ("
              << lo[0] << "," << lo[1] << "," << lo[2] << ")-(" << hi[0] << "," << hi[1] << "," << hi[2] << ")" << std::endl;
    std::vector<float> result;
    if (gpu_available()) {
      select_gpu(bs, o);
      // default: in place when the files' boxes are (nearly) disjoint, else redistribute the union
      // (--in-place / --redistribute / --ring force a strategy)
      const int mode = o.mode >= 0 ? o.mode : B200KNN_MODE_AUTO;
      result = run_gpu(bs, pts, 0, o, mode == B200KNN_MODE_REDISTRIBUTE ? B200KNN_MODE_AUTO : mode);
    } else {
      result = run_cpu(bs, pts, o);
    }
    if (bs.rank() == 0) std::cout << "done all queries..." << std::endl;
    write_at(per_rank_name(o.output, bs.rank()), result.data(), (int64_t)result.size(), 0, true, (int64_t)result.size());
    bs.barrier();
    return 0;
  } catch (const std::exception& e) {
    std::cerr << "fatal: " << e.what() << std::endl;
    return 2;
  }
}
