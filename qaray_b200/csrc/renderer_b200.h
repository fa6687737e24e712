// renderer_b200.h — C++ host side of the drop-in boundary: the reference's renderer interface
// (RendererParam, Renderer::Init / ComputeScene / Render / Terminate; src/renderers/renderer.h:47-110, driven from
// src/main.cpp:46-59) over the C ABI of include/qaray_b200.h. Header-only; links against libqaray_b200.so.
// Everything that computes happens behind qg_*: there is no CPU path here.
//
// Error behaviour follows the reference: LoadScene returns 0/1 and prints (src/parser/xmlload.cpp:71-95); the
// other calls print the library's message and return false instead of throwing (the reference uses no exceptions).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "../../include/qaray_b200.h"

namespace qaray_b200 {

// RendererParam (renderer.h:47-68), photon-map switches included: ComputeScene builds the maps (renderer.cpp:118-291).
struct RendererParam {
  bool useSRGB = true;
  size_t sppMax = 8;
  size_t sppMin = 4;
  bool usePhotonMap = false;
  size_t photonMapSize = 10000, causticsMapSize = 1000;
  size_t photonMapBounce = 20, causticsMapBounce = 20;
  float photonMapRadius = 0.2f, causticsMapRadius = 1.0f;
  int maxBounce = 5;  // Material::maxBounce (src/core/material.cpp:31)
  unsigned seed = 0x5EEDu;
  int width = 0, height = 0;  // 0 = the XML camera's
  // which class `MtlBlinn` names: a compile-time alias in the reference (src/materials/materials.h:57-61), a run-time
  // switch here (QG_CLASS_*); default = MtlBlinn_PhotonMap like the reference's build
  int materialClass = QG_CLASS_PHOTONMAP;
  bool sortByMaterial = false;  // shade each depth's hits in material order (qgRenderParams.sort_by_material)
  void SetPhotonMappingFlag(bool f) { usePhotonMap = f; }
  void SetPhotonMapSize(size_t s) { photonMapSize = s; }
  void SetCausticsMapSize(size_t s) { causticsMapSize = s; }
  void SetSPPMax(int spp) { sppMax = (size_t) spp; }
  void SetSPPMin(int spp) { sppMin = (size_t) spp; }
  void SetSRGBFlag(bool f) { useSRGB = f; }
};

class Renderer_B200 {
 protected:
  RendererParam &param;
  qgScene *scene = nullptr;
  qgMulti *multi = nullptr;  // -gpus N: one scene replica, host thread and NCCL rank per device (Renderer_MPI's rank model)
  int numGpus = 1, tileSplit = 1;
  std::string assetDir, rankImagePrefix;
  bool saveRankImages = false;
  size_t pixelW = 0, pixelH = 0;
  size_t mpiSize = 1, mpiRank = 0;  // Renderer::mpiSize / mpiRank: the interleaved 4x4-tile shard (renderer.cpp:383-387)
 public:
  std::vector<uint8_t> colorBuffer, sampleCountBuffer, maskBuffer;  // FrameBuffer planes (src/fb/framebuffer.h:35-95)
  std::vector<float> depthBuffer, linearBuffer;
  qgRenderStats stats{};
  qgPhotonStats photonStats{};
  bool mapsBuilt = false;

  explicit Renderer_B200(RendererParam &p) : param(p) {}
  ~Renderer_B200() { Terminate(); }
  void SetShard(size_t rank, size_t size) { mpiRank = rank; mpiSize = size; }
  void SetAssetDir(const std::string &d) { assetDir = d; }
  // Renderer_MPI (src/renderers/Renderer_MPI.cpp:33-55) inside one process: G = gpus ranks, tileSplit = G_t of them share
  // a sample range by the interleaved-tile rule, G / G_t sample ranges. rankImages: also write rank_<r>_*Buffer.png (:132-138)
  void SetGpus(int gpus, int tile_split = 1, bool rankImages = false) { numGpus = gpus; tileSplit = tile_split; saveRankImages = rankImages; }
  void SetRankImagePrefix(const std::string &p) { rankImagePrefix = p; }
  size_t W() const { return pixelW; }
  size_t H() const { return pixelH; }

  bool Init(int device = 0)
  {
    if (qg_init(device) != QG_OK) { fprintf(stderr, "Renderer_B200::Init: %s\n", qg_last_error()); return false; }
    if (numGpus > 1) {  // devices device .. device + G - 1
      std::vector<int32_t> devs((size_t) numGpus);
      for (int g = 0; g < numGpus; ++g) devs[(size_t) g] = device + g;
      if (qg_multi_create(devs.data(), numGpus, &multi) != QG_OK) { fprintf(stderr, "Renderer_B200::Init: %s\n", qg_last_error()); return false; }
    }
    return true;
  }
  // LoadScene(filename): 1 on success, 0 on failure
  int LoadScene(const char *filename)
  {
    if (multi) {
      if (qg_multi_load_xml(multi, filename, assetDir.c_str()) != QG_OK) {
        printf("Failed to load the file \"%s\": %s\n", filename, qg_last_error());
        return 0;
      }
      scene = qg_multi_scene(multi, 0);  // owned by multi
      return 1;
    }
    if (scene) { qg_scene_destroy(scene); scene = nullptr; }
    if (qg_scene_load_xml(filename, assetDir.c_str(), &scene) != QG_OK) {
      printf("Failed to load the file \"%s\": %s\n", filename, qg_last_error());
      return 0;
    }
    return 1;
  }
  // Renderer::ComputeScene: image size and plane allocation (the camera frame itself is computed by the library)
  bool ComputeScene()
  {
    qgSceneInfo info;
    if (!scene || qg_scene_info(scene, &info) != QG_OK) return false;
    pixelW = (size_t) (param.width > 0 ? param.width : info.camera.width);
    pixelH = (size_t) (param.height > 0 ? param.height : info.camera.height);
    const size_t n = pixelW * pixelH;
    colorBuffer.assign(3 * n, 0);
    depthBuffer.assign(n, 0.f);
    sampleCountBuffer.assign(n, 0);
    maskBuffer.assign(n, 0);
    linearBuffer.assign(3 * n, 0.f);
    // Initialize Photon Map (renderer.cpp:115-291). The reference leaves "TODO: Photon Map for MPI Mode"; here every GPU of
    // -gpus N builds the maps on its own scene replica from the same seed — emission e draws from the stream of (e, map, seed),
    // so the replicas hold identical maps and the shards gather from the same photons
    if (param.usePhotonMap && param.photonMapSize > 0 && param.causticsMapSize > 0) {
      qgPhotonParams pp{};
      pp.photon_map_size = (int32_t) param.photonMapSize; pp.photon_map_bounce = (int32_t) param.photonMapBounce;
      pp.photon_map_radius = param.photonMapRadius;
      pp.caustics_map_size = (int32_t) param.causticsMapSize; pp.caustics_map_bounce = (int32_t) param.causticsMapBounce;
      pp.caustics_map_radius = param.causticsMapRadius;
      pp.seed = param.seed;
      for (int g = 0; g < (multi ? numGpus : 1); ++g) {
        qgScene *replica = multi ? qg_multi_scene(multi, g) : scene;
        if (qg_photon_maps_build(replica, &pp, g == 0 ? &photonStats : nullptr) != QG_OK) {
          fprintf(stderr, "Renderer_B200::ComputeScene: %s\n", qg_last_error());
          return false;
        }
      }
      printf("\nPhoton Map: %llu photons, Caustics Map: %llu photons, %.3f ms\n", (unsigned long long) photonStats.photons[0],
             (unsigned long long) photonStats.photons[1], photonStats.device_ms);
      mapsBuilt = true;
    }
    return true;
  }
  // Renderer::Render (batch): ThreadRender over this shard's tiles; fills the planes PixelRender fills
  bool Render()
  {
    qgRenderParams rp{};
    rp.spp = (int32_t) param.sppMax;
    rp.max_bounce = param.maxBounce;
    rp.srgb = param.useSRGB ? 1 : 0;
    rp.seed = param.seed;
    rp.width = (int32_t) pixelW;
    rp.height = (int32_t) pixelH;
    rp.tile_rank = (int32_t) mpiRank;
    rp.tile_ranks = (int32_t) mpiSize;
    rp.material_class = param.materialClass;
    rp.sort_by_material = param.sortByMaterial ? 1 : 0;
    rp.spp_min = param.sppMin < param.sppMax ? (int32_t) param.sppMin : 0;  // SuperSamplerHalton's adaptive stop
    rp.use_photon_map = mapsBuilt ? 1 : 0;  // scene->usePhotonMap = param.usePhotonMap (renderer.cpp:117)
    if (multi) {
      if (mpiSize > 1) { fprintf(stderr, "Renderer_B200: -shard and -gpus do not combine\n"); return false; }
      rp.tile_rank = 0;
      rp.tile_ranks = 1;
      const std::string pre = rankImagePrefix;
      if (qg_multi_render(multi, &rp, tileSplit, saveRankImages ? pre.c_str() : nullptr, linearBuffer.data(), colorBuffer.data(), depthBuffer.data(),
                          sampleCountBuffer.data(), maskBuffer.data(), &stats) != QG_OK) {
        fprintf(stderr, "Renderer_B200::Render: %s\n", qg_last_error());
        return false;
      }
      return true;
    }
    if (qg_render(scene, &rp, linearBuffer.data(), colorBuffer.data(), depthBuffer.data(), sampleCountBuffer.data(), maskBuffer.data(),
                  &stats) != QG_OK) {
      fprintf(stderr, "Renderer_B200::Render: %s\n", qg_last_error());
      return false;
    }
    return true;
  }
  // the PNG set of Renderer_MPI::Render (src/renderers/Renderer_MPI.cpp:134-138,188-193): FrameBuffer::SaveImage, SaveZImage
  // (depth normalised, framebuffer.cpp:62-86,113-135) and SaveSampleCountImage (sample counts stretched to 0..255, :88-107)
  bool SaveImages(const std::string &prefix = "", const std::string &suffix = "")
  {
    const int w = (int) pixelW, h = (int) pixelH;
    return qg_save_png_rgb8((prefix + "colorBuffer" + suffix + ".png").c_str(), colorBuffer.data(), w, h) == QG_OK &&
           qg_save_png_depth((prefix + "depthBuffer" + suffix + ".png").c_str(), depthBuffer.data(), w, h) == QG_OK &&
           qg_save_png_sample_count((prefix + "sampleBuffer" + suffix + ".png").c_str(), sampleCountBuffer.data(), w, h) == QG_OK;
  }
  void Terminate()
  {
    if (multi) { qg_multi_destroy(multi); multi = nullptr; scene = nullptr; }
    if (scene) { qg_scene_destroy(scene); scene = nullptr; }
  }
};

}  // namespace qaray_b200
