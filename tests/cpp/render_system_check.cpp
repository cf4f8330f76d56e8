// TEST CODE: compiles include/cavept_render_system.hpp as C++20 with -Wall -Wextra -Werror and drives it.
//   without a GPU: Initialize must fail with CAVEPT_ERR_CUDA (no CPU path) and every call must stay an error code;
//   with a GPU (argv[1] == "gpu"): the reference's call sequence -- SetPathTracerMode(INTERACTIVE),
//   RequestPathTracerUpdate, Dispatch x3 with a dirty camera in between, read-back -- on a two-triangle scene.
#include <cstdio>
#include <cstring>

#include "cavept_render_system.hpp"

using namespace cavept_rs;

#define CHECK(cond)                                                      \
    do {                                                                 \
        if (!(cond)) {                                                   \
            std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
            return 1;                                                    \
        }                                                                \
    } while (0)

int main(int argc, char** argv) {
    const bool gpu = argc > 1 && std::strcmp(argv[1], "gpu") == 0;
    // a floor quad facing +y and an emissive quad above it facing -y
    const float pos[] = { -2, 0, -2, -2, 0, 2, 2, 0, 2, 2, 0, -2, -1, 2, -1, 1, 2, -1, 1, 2, 1, -1, 2, 1 };
    const float nrm[] = { 0, 1, 0, 0, 1, 0, 0, 1, 0, 0, 1, 0, 0, -1, 0, 0, -1, 0, 0, -1, 0, 0, -1, 0 };
    const uint32_t idx_floor[] = { 0, 1, 2, 0, 2, 3 };
    const uint32_t idx_lamp[] = { 0, 1, 2, 0, 2, 3 };
    cavept_mesh meshes[2] = { { pos, nrm, idx_floor, 4, 6 }, { pos + 12, nrm + 12, idx_lamp, 4, 6 } };
    cavept_instance inst[2] = {};
    for (int k = 0; k < 2; ++k) {
        inst[k].mesh = k;
        inst[k].material = k;
        for (int d = 0; d < 4; ++d) inst[k].transform[5 * d] = 1.0f;
    }
    cavept_material mats[2] = { { { 0.7f, 0.7f, 0.7f }, 0.9f, 0.0f, 0.0f }, { { 1.0f, 1.0f, 1.0f }, 1.0f, 0.0f, 5.0f } };
    SceneView scene;
    scene.meshes = meshes;
    scene.mesh_count = 2;
    scene.instances = inst;
    scene.instance_count = 2;
    scene.materials = mats;
    scene.material_count = 2;
    CameraView cam = { { 0.0f, 1.0f, -3.0f }, { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 }, 50.0f };

    CHECK(!IsPathTracerActive());
    RenderSystem& rs = Instance();
    const int rc = rs.Initialize(64, 36, 5);
    if (!gpu) {
        CHECK(rc == CAVEPT_ERR_CUDA || rc == CAVEPT_OK);  // CAVEPT_OK only if this host happens to have a GPU
        if (rc != CAVEPT_OK) {
            CHECK(rs.RequestPathTracerUpdate(cam, scene) == CAVEPT_ERR_NOT_READY);
            CHECK(rs.Dispatch() == CAVEPT_ERR_NOT_READY);
            CHECK(!IsPathTracerActive());
            std::printf("OK (no device: every call is an error code)\n");
            return 0;
        }
    }
    CHECK(rc == CAVEPT_OK);
    RequestPathTracerUpdate(cam, scene);  // mode NONE: PathTracer::Update returns at once (path_tracer.cpp:51-52)
    CHECK(!IsPathTracerActive());
    SetPathTracerMode(PathTracerMode::INTERACTIVE);
    CHECK(!IsPathTracerActive());  // no BVH resident yet (path_tracer.cpp:317-319)
    RequestPathTracerUpdate(cam, scene);
    CHECK(IsPathTracerActive());
    CHECK(rs.Dispatch() == CAVEPT_OK);
    CHECK(rs.Dispatch() == CAVEPT_OK);
    std::vector<float> acc;
    CHECK(rs.ReadAccumulator(acc) == CAVEPT_OK);
    CHECK(acc.size() == 64u * 36u * 4u && acc[3] == 2.0f);
    scene.dirty_flags = SCENE_DIRTY_CAMERA;  // the next frame overwrites (c_sceneDirty)
    cam.position[1] = 1.2f;
    RequestPathTracerUpdate(cam, scene);
    CHECK(rs.Dispatch() == CAVEPT_OK);
    CHECK(rs.ReadAccumulator(acc) == CAVEPT_OK && acc[3] == 1.0f);
    std::vector<uint8_t> tone;
    CHECK(rs.ReadToneMappedRgba8(tone) == CAVEPT_OK && tone[3] == 255);
    float lit = 0.0f;
    for (size_t i = 0; i < acc.size(); i += 4) lit += acc[i];
    CHECK(lit > 1.0f);  // the lamp is seen
    scene.dirty_flags = SCENE_DIRTY_WORLD;
    inst[1].transform[13] = 0.5f;
    RequestPathTracerUpdate(cam, scene);
    CHECK(rs.Dispatch() == CAVEPT_OK);
    SetPathTracerMode(PathTracerMode::NONE);
    CHECK(!IsPathTracerActive() && rs.Dispatch() == CAVEPT_ERR_NOT_READY);
    CHECK(rs.SetPathTracerMode(static_cast<PathTracerMode>(7)) == CAVEPT_ERR_INVALID_ARG);  // CRASH_NOW_MSG("Invalid mode") in the reference
    rs.Shutdown();
    std::printf("OK (device path)\n");
    return 0;
}
