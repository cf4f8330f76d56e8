// cavept_render_system.hpp -- header-only C++ facade over include/cavept.h with the SHAPE of CaveEngine's
// renderer-facing path-tracer surface, ECS-free so that it compiles (and is compile-checked by
// tests/test_render_system_header.py) outside the engine:
//
//   cave/engine/renderer/path_tracer_render_system.h:10-21
//       enum class PathTracerMode { NONE, INTERACTIVE, TILED };
//       void RequestPathTracerUpdate(const CameraComponent&, Scene&);
//       void SetPathTracerMode(PathTracerMode);
//       bool IsPathTracerActive();
//       void BindPathTracerData(IGraphicsManager&);   void UnbindPathTracerData(IGraphicsManager&);
//   backed by ONE file-static PathTracer (path_tracer_render_system.cpp:8-10) whose Update()
//   (renderer/path_tracer/path_tracer.cpp:49-70) uploads once and then re-uploads instances / materials when
//   Scene::GetDirtyFlags() says so (:156-158, :197-199).
//
// Here the engine's CameraComponent / Scene are replaced by two plain views (CameraView, SceneView) that an engine
// adapter fills from its ECS (INTEGRATION.md section 2 shows that adapter); everything else keeps the reference's
// names, argument meaning and error behaviour (an invalid mode is CRASH_NOW_MSG there, an exception-free error code
// here; Update() is a no-op in mode NONE).  Bind/Unbind are no-ops: the buffers the reference binds to shader slots
// (path_tracer.cpp:324-340) live inside the cavept context, and the pass itself is Dispatch() below, the
// counterpart of PathTracerPassFunc (render_graph/common_passes.cpp:977-999).
#ifndef CAVEPT_RENDER_SYSTEM_HPP_
#define CAVEPT_RENDER_SYSTEM_HPP_

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "cavept.h"

namespace cavept_rs {

enum class PathTracerMode {  // path_tracer_render_system.h:10-14; values equal CAVEPT_MODE_*
    NONE,
    INTERACTIVE,
    TILED,
};
static_assert(static_cast<int>(PathTracerMode::NONE) == CAVEPT_MODE_NONE && static_cast<int>(PathTracerMode::INTERACTIVE) == CAVEPT_MODE_INTERACTIVE &&
                  static_cast<int>(PathTracerMode::TILED) == CAVEPT_MODE_TILED,
              "enum values follow the reference");

// Scene::GetDirtyFlags(), cave/engine/scene/scene.h:74-79
enum SceneDirtyFlags : uint32_t {
    SCENE_DIRTY_NONE = 1u << 0,
    SCENE_DIRTY_WORLD = 1u << 1,
    SCENE_DIRTY_CAMERA = 1u << 2,
    SCENE_DIRTY_LIGHT = 1u << 3,
};

// What RequestPathTracerUpdate reads from CameraComponent (render_system.cpp:89-93 -> c_camera*, cbuffer.hlsl.h:193-207)
struct CameraView {
    float position[3];
    float right[3];
    float up[3];
    float forward[3];
    float fov_degree;
};

// What PathTracer::Update reads from Scene: MeshRendererComponent + TransformComponent + MaterialComponent + MeshAsset
// (path_tracer.cpp:203-310, :120-201), flattened by the caller.  Pointers must stay valid during the call only.
struct SceneView {
    const cavept_mesh* meshes = nullptr;
    int mesh_count = 0;
    const cavept_instance* instances = nullptr;
    int instance_count = 0;
    const cavept_material* materials = nullptr;
    int material_count = 0;
    uint32_t dirty_flags = SCENE_DIRTY_NONE;
};

// The file-static PathTracer of path_tracer_render_system.cpp:8-10.
class RenderSystem {
public:
    RenderSystem() = default;
    RenderSystem(const RenderSystem&) = delete;
    RenderSystem& operator=(const RenderSystem&) = delete;
    ~RenderSystem() { Shutdown(); }

    // Creates the context (the reference constructs its PathTracer statically; a GPU context needs a size).
    int Initialize(int p_width, int p_height, int p_max_bounce = 5, int p_device = -1, int p_bvh_mode = CAVEPT_BVH_PARITY) {
        Shutdown();
        cavept_config cfg{};
        cfg.device = p_device;
        cfg.width = p_width;
        cfg.height = p_height;
        cfg.max_bounce = p_max_bounce;
        cfg.trace_mode = CAVEPT_TRACE_FAST;
        cfg.bvh_mode = p_bvh_mode;
        const int rc = cavept_create(&cfg, &m_ctx);
        if (rc == CAVEPT_OK) cavept_set_mode(m_ctx, static_cast<int>(m_mode));
        return rc;
    }
    void Shutdown() {
        if (m_ctx) cavept_destroy(m_ctx);
        m_ctx = nullptr;
        m_uploaded = false;
        m_frameIndex = 0;
    }

    // SetPathTracerMode, path_tracer_render_system.cpp:17-19
    int SetPathTracerMode(PathTracerMode p_mode) {
        m_mode = p_mode;
        return m_ctx ? cavept_set_mode(m_ctx, static_cast<int>(p_mode)) : CAVEPT_OK;
    }
    // IsPathTracerActive, :21-23 -> PathTracer::IsActive (path_tracer.cpp:312-322)
    bool IsPathTracerActive() const { return m_ctx && cavept_is_active(m_ctx) != 0; }

    // RequestPathTracerUpdate, :12-15 -> PathTracer::Update (path_tracer.cpp:49-70): nothing in mode NONE; the
    // first call uploads the scene and builds the BVHs (CreateAccelStructure), later calls re-upload what the dirty
    // flags name (UpdateAccelStructure :156-158 materials on any dirty flag, :197-199 instances on SCENE_DIRTY_WORLD).
    int RequestPathTracerUpdate(const CameraView& p_camera, const SceneView& p_scene) {
        if (!m_ctx) return CAVEPT_ERR_NOT_READY;
        if (m_mode == PathTracerMode::NONE) return CAVEPT_OK;
        int rc = CAVEPT_OK;
        bool dirty = false;
        if (!m_uploaded) {
            rc = cavept_upload_scene(m_ctx, p_scene.meshes, p_scene.mesh_count, p_scene.instances, p_scene.instance_count, p_scene.materials, p_scene.material_count);
            if (rc != CAVEPT_OK) return rc;
            m_uploaded = true;
            dirty = true;
        } else {
            if (p_scene.dirty_flags & SCENE_DIRTY_WORLD) {
                rc = cavept_update_instances(m_ctx, p_scene.instances, p_scene.instance_count);
                if (rc != CAVEPT_OK) return rc;
            }
            if (p_scene.dirty_flags != SCENE_DIRTY_NONE) {
                rc = cavept_update_materials(m_ctx, p_scene.materials, p_scene.material_count);
                if (rc != CAVEPT_OK) return rc;
                dirty = true;
            }
        }
        if (dirty || (p_scene.dirty_flags & SCENE_DIRTY_CAMERA) || !m_cameraSet) {
            rc = cavept_set_camera(m_ctx, p_camera.position, p_camera.right, p_camera.up, p_camera.forward, p_camera.fov_degree);  // marks c_sceneDirty
            m_cameraSet = rc == CAVEPT_OK;
        }
        return rc;
    }

    // The reference binds five structured buffers to shader slots around the dispatch; the context owns them here.
    void BindPathTracerData() const {}
    void UnbindPathTracerData() const {}

    // PathTracerPassFunc's Dispatch (common_passes.cpp:990-997): one frame with the running c_frameIndex
    // (render_system.cpp:102), asynchronous on the context's stream.
    int Dispatch() {
        if (!IsPathTracerActive()) return CAVEPT_ERR_NOT_READY;
        BindPathTracerData();
        const int rc = cavept_render(m_ctx, m_frameIndex, 1);
        UnbindPathTracerData();
        if (rc == CAVEPT_OK) ++m_frameIndex;
        return rc;
    }

    // RG_RES_PATHTRACER read-back (the reference leaves it on the GPU; the UI samples the tone-mapped image).
    int ReadAccumulator(std::vector<float>& p_out) {
        if (!m_ctx) return CAVEPT_ERR_NOT_READY;
        p_out.resize(static_cast<size_t>(m_width()) * static_cast<size_t>(m_height()) * 4u);
        return cavept_read_accum(m_ctx, p_out.data(), p_out.size() * sizeof(float));
    }
    int ReadToneMappedRgba8(std::vector<uint8_t>& p_out) {
        if (!m_ctx) return CAVEPT_ERR_NOT_READY;
        p_out.resize(static_cast<size_t>(m_width()) * static_cast<size_t>(m_height()) * 4u);
        return cavept_read_tonemapped_rgba8(m_ctx, p_out.data(), p_out.size());
    }

    std::string LastError() const { return m_ctx ? std::string(cavept_last_error(m_ctx)) : std::string("no context"); }
    cavept_ctx* Context() const { return m_ctx; }
    int FrameIndex() const { return m_frameIndex; }

private:
    int m_width() const {
        int w = 0;
        cavept_get_image_size(m_ctx, &w, nullptr);
        return w;
    }
    int m_height() const {
        int h = 0;
        cavept_get_image_size(m_ctx, nullptr, &h);
        return h;
    }
    cavept_ctx* m_ctx = nullptr;
    PathTracerMode m_mode = PathTracerMode::NONE;
    bool m_uploaded = false;
    bool m_cameraSet = false;
    int m_frameIndex = 0;
};

// ---- the five free functions of path_tracer_render_system.h over one process-wide instance ------------------
inline RenderSystem& Instance() {
    static RenderSystem s_glob;  // path_tracer_render_system.cpp:8-10
    return s_glob;
}
inline void RequestPathTracerUpdate(const CameraView& p_camera, const SceneView& p_scene) { (void)Instance().RequestPathTracerUpdate(p_camera, p_scene); }
inline void SetPathTracerMode(PathTracerMode p_mode) { (void)Instance().SetPathTracerMode(p_mode); }
inline bool IsPathTracerActive() { return Instance().IsPathTracerActive(); }
inline void BindPathTracerData() { Instance().BindPathTracerData(); }
inline void UnbindPathTracerData() { Instance().UnbindPathTracerData(); }

}  // namespace cavept_rs

#endif  // CAVEPT_RENDER_SYSTEM_HPP_
