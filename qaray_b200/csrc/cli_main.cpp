// cli_main.cpp — main.cpp-style driver for the B200 path: same flags as the reference's CLI
// (src/main.cpp:17-44: -spp N, -sppMin N, -sppMax N, -bounce N, -srgb 0|1, <scene.xml>; -batch is implied, there is
// no GUI), plus -assets DIR (where the XML's relative .obj/.png names resolve), -size W H, -seed S, -device D,
// -shard RANK SIZE (Renderer::mpiRank / mpiSize), -gpus N [-tilesplit T] [-rank-images] (Renderer_MPI over NCCL in one process), -mtl CLASS (which class MtlBlinn names) and -out PREFIX. Writes colorBuffer.png / depthBuffer.png /
// sampleBuffer.png like Renderer_MPI::Render and prints one line of ray statistics.
#include <cstring>
#include <string>

#include "renderer_b200.h"

using namespace qaray_b200;

int main(int argc, char **argv)
{
  RendererParam param;
  const char *file = nullptr;
  std::string assets, out;
  int device = 0, rank = 0, size = 1, gpus = 1, tile_split = 1;
  bool rank_images = false;
  if (argc < 2) {
    fprintf(stderr, "Error: insufficient input\n");
    return -1;
  }
  for (int i = 1; i < argc; ++i) {
    const std::string s(argv[i]);
    auto next = [&]() -> const char * { return i + 1 < argc ? argv[++i] : "0"; };
    if (s == "-batch") {
    } else if (s == "-spp") {
      const int v = atoi(next());
      param.SetSPPMax(v);
      param.SetSPPMin(v);
    } else if (s == "-sppMin") param.SetSPPMin(atoi(next()));
    else if (s == "-sppMax") param.SetSPPMax(atoi(next()));
    else if (s == "-bounce") param.maxBounce = atoi(next());
    else if (s == "-srgb") param.SetSRGBFlag(atoi(next()) != 0);
    else if (s == "-threads") next();  // the TBB/OpenMP thread count has no meaning here
    else if (s == "-use-photon-map") param.SetPhotonMappingFlag(true);
    else if (s == "-photon-map-size") param.SetPhotonMapSize((size_t) atoi(next()));
    else if (s == "-caustics-map-size") param.SetCausticsMapSize((size_t) atoi(next()));
    else if (s == "-assets") assets = next();
    else if (s == "-out") out = next();
    else if (s == "-seed") param.seed = (unsigned) strtoul(next(), nullptr, 0);
    else if (s == "-device") device = atoi(next());
    else if (s == "-size") { param.width = atoi(next()); param.height = atoi(next()); }
    else if (s == "-shard") { rank = atoi(next()); size = atoi(next()); }
    else if (s == "-gpus") gpus = atoi(next());          // one NCCL rank per device device .. device + N - 1 (Renderer_MPI's rank model)
    else if (s == "-tilesplit") tile_split = atoi(next());  // G_t of the G = G_t x G_s split; default 1 = sample shards
    else if (s == "-rank-images") rank_images = true;      // also rank_<r>_{color,depth,sample}Buffer.png (Renderer_MPI.cpp:132-138)
    else if (s == "-sortmtl") param.sortByMaterial = true;
    else if (s == "-mtl") {  // the reference's compile-time `using MtlBlinn = ...` (materials.h:57-61)
      const std::string c(next());
      if (c == "photonmap") param.materialClass = QG_CLASS_PHOTONMAP;
      else if (c == "pathtracing") param.materialClass = QG_CLASS_PATHTRACING;
      else if (c == "montecarlogi") param.materialClass = QG_CLASS_MONTECARLOGI;
      else if (c == "blinn_basic") param.materialClass = QG_CLASS_BLINN_BASIC;
      else if (c == "phong_basic") param.materialClass = QG_CLASS_PHONG_BASIC;
      else { fprintf(stderr, "Error: unknown material class %s\n", c.c_str()); return -1; }
    }
    else file = argv[i];
  }
  if (!file) {
    fprintf(stderr, "Error: no scene file\n");
    return -1;
  }
  Renderer_B200 renderer(param);
  renderer.SetAssetDir(assets);
  renderer.SetShard((size_t) rank, (size_t) size);
  if (gpus > 1) {
    renderer.SetGpus(gpus, tile_split, rank_images);
    renderer.SetRankImagePrefix(out);
  }
  if (!renderer.Init(device)) return 1;
  if (!renderer.LoadScene(file)) return 1;
  if (!renderer.ComputeScene()) return 1;
  if (!renderer.Render()) return 1;
  const qgRenderStats &st = renderer.stats;
  const double rays = (double) st.closest_rays + (double) st.shadow_rays;
  printf("Elapsed Time: %.3f ms  (%zux%zu, %llu samples, %llu closest + %llu shadow rays, %.1f Mrays/s, %llu launches)\n", st.device_ms,
         renderer.W(), renderer.H(), (unsigned long long) st.samples, (unsigned long long) st.closest_rays,
         (unsigned long long) st.shadow_rays, st.device_ms > 0 ? rays / st.device_ms * 1e-3 : 0.0, (unsigned long long) st.kernel_launches);
  // Renderer_MPI::Render names the composited planes *_MPI.png (Renderer_MPI.cpp:188-193)
  if (!renderer.SaveImages(out, gpus > 1 ? "_MPI" : "")) {
    fprintf(stderr, "cannot write images: %s\n", qg_last_error());
    return 1;
  }
  renderer.Terminate();
  qg_shutdown();
  return 0;
}
