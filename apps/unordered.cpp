// cudaMpiKNN_unorderedData — one flat float3 file in arbitrary order, one output file of N floats.
// Same command line, input and output as the reference program of the same name
// (/root/reference/unorderedDataVariant.cu:105-238):
//   knnrun -n <ranks> ./cudaMpiKNN_unorderedData points.float3 -o distances.float -k 100 [-r R] [-g G]
// Default strategy: scatter points to spatial owners, bbox-culled exchange, results returned to the
// reading rank (REDISTRIBUTE); --ring selects the all-pairs fused ring, --in-place the culled
// exchange without redistribution.
#include "knn_apps.hpp"

using namespace knnapp;

int main(int ac, char** av) {
  try {
    const Options o = parse_args(ac, av, "input file name", "output file name");
    Bootstrap bs;
    int64_t begin = 0, total = 0;
    std::vector<float> result;
    if (gpu_available()) {
      total = count_points(o.input);
      begin = total * bs.rank() / bs.size();
      const int64_t n = total * (bs.rank() + 1) / bs.size() - begin;
      std::cout << "#" << bs.rank() << "/" << bs.size() << ": got " << n << " points to work on" << std::endl;
      select_gpu(bs, o);
      const int mode = o.mode >= 0 ? o.mode : B200KNN_MODE_REDISTRIBUTE;
      // file -> device -> file; the results never exist as one host array
      run_gpu_from_file(bs, o.input, begin, n, o, mode, o.output, total);
      return 0;
    } else {
      std::vector<float> pts = read_points(o.input, bs.rank(), bs.size(), &begin, &total);
      std::cout << "#" << bs.rank() << "/" << bs.size() << ": got " << pts.size() / 3 << " points to work on" << std::endl;
      result = run_cpu(bs, pts, o);
    }
    if (bs.rank() == 0) std::cout << "done all queries..." << std::endl;
    // rank 0 creates the file at its final size, then every rank writes its slice at its offset:
    // the same bytes as the reference's rank-ordered append (unorderedDataVariant.cu:229-237)
    if (bs.rank() == 0) write_at(o.output, result.data(), (int64_t)result.size(), begin, true, total);
    bs.barrier();
    if (bs.rank() != 0) write_at(o.output, result.data(), (int64_t)result.size(), begin, false, total);
    bs.barrier();
    return 0;
  } catch (const std::exception& e) {
    std::cerr << "fatal: " << e.what() << std::endl;
    return 2;
  }
}
