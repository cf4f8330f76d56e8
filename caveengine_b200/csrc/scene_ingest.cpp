// Asset ingest for the path-tracer pass: the engine's binary
// `.mesh` files and YAML `.scene` files -> the flat arrays cavept_upload_scene takes.
// Host-side file parsing only (the reference does this on the CPU too); every byte of
// geometry still goes through the device builders and kernels.
//
// Reference behaviour restated here (paths relative to the CaveEngine repository root):
//   * MeshAsset::SerializeBinary, cave/engine/assets/mesh_asset.cpp:45-61, over Archive's
//     "u64 count + raw elements" vectors, cave/engine/core/io/archive.h:56-75
//   * the scene YAML written by Scene serialisation (cave/engine/scene/scene.cpp:370,437):
//     `entities:` list of {id, NameComponent, HierarchyComponent{parent_id},
//     TransformComponent{translation, rotation(xyzw), scale}, MeshRendererComponent{mesh_id,
//     materials}, MaterialComponent{base_color, metallic, roughness, emissive}}
//   * TransformComponent::GetLocalMatrix, cave/engine/scene/transform_component.cpp:16-21
//     (translate * toMat4(quat) * scale, glm fp32 operation order)
//   * UpdateHierarchy, cave/engine/systems/ecs_systems.cpp:139-171 (world = parent locals
//     multiplied on from the left, one ancestor at a time)
//   * PathTracer::CreateAccelStructure / UpdateAccelStructure,
//     cave/engine/renderer/path_tracer/path_tracer.cpp:120-252: one instance per entity with a
//     MeshRendererComponent and a TransformComponent, material = materials[0] of the renderer
//     (`is_visible` is NOT consulted by the path tracer).
//   * asset lookup by GUID through `<file>.meta` side-cars (`guid:` key).
#include <algorithm>
#include <cerrno>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include <dirent.h>
#include <sys/stat.h>

#include "../../include/cavept.h"

namespace {

// ------------------------------------------------------------------------------------------
// glm fp32 restatements (thirdparty/glm, default packed_highp: no SIMD, no FMA)
// ------------------------------------------------------------------------------------------
struct Mat4 {
    float m[16];  // column-major, m[c*4 + r]
};

// operator*(mat4, mat4): thirdparty/glm/glm/detail/type_mat4x4.inl:660-698 (tmp = A0*b.x; += A1*b.y; ...)
Mat4 mat_mul(const Mat4& a, const Mat4& b) {
    Mat4 out;
    for (int c = 0; c < 4; ++c) {
        for (int r = 0; r < 4; ++r) {
            float t = a.m[0 * 4 + r] * b.m[c * 4 + 0];
            t += a.m[1 * 4 + r] * b.m[c * 4 + 1];
            t += a.m[2 * 4 + r] * b.m[c * 4 + 2];
            t += a.m[3 * 4 + r] * b.m[c * 4 + 3];
            out.m[c * 4 + r] = t;
        }
    }
    return out;
}

Mat4 mat_identity() {
    Mat4 i{};
    i.m[0] = i.m[5] = i.m[10] = i.m[15] = 1.0f;
    return i;
}

// glm::translate(mat4(1), v): Result[3] = m[0]*v[0] + m[1]*v[1] + m[2]*v[2] + m[3]
Mat4 mat_translate(const float v[3]) {
    const Mat4 i = mat_identity();
    Mat4 out = i;
    for (int r = 0; r < 4; ++r) out.m[12 + r] = ((i.m[0 + r] * v[0] + i.m[4 + r] * v[1]) + i.m[8 + r] * v[2]) + i.m[12 + r];
    return out;
}

// glm::scale(mat4(1), v): Result[k] = m[k] * v[k]
Mat4 mat_scale(const float v[3]) {
    const Mat4 i = mat_identity();
    Mat4 out = i;
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 4; ++r) out.m[c * 4 + r] = i.m[c * 4 + r] * v[c];
    return out;
}

// glm::toMat4(quat) = mat4(mat3_cast(q)), thirdparty/glm/glm/gtc/quaternion.inl:47-72
Mat4 mat_from_quat(float x, float y, float z, float w) {
    const float qxx = x * x, qyy = y * y, qzz = z * z, qxz = x * z, qxy = x * y, qyz = y * z, qwx = w * x, qwy = w * y, qwz = w * z;
    Mat4 out = mat_identity();
    out.m[0] = 1.0f - 2.0f * (qyy + qzz);
    out.m[1] = 2.0f * (qxy + qwz);
    out.m[2] = 2.0f * (qxz - qwy);
    out.m[4] = 2.0f * (qxy - qwz);
    out.m[5] = 1.0f - 2.0f * (qxx + qzz);
    out.m[6] = 2.0f * (qyz + qwx);
    out.m[8] = 2.0f * (qxz + qwy);
    out.m[9] = 2.0f * (qyz - qwx);
    out.m[10] = 1.0f - 2.0f * (qxx + qyy);
    return out;
}

// TransformComponent::GetLocalMatrix: translationMatrix * rotationMatrix * scaleMatrix
Mat4 local_matrix(const float t[3], const float q[4], const float s[3]) {
    return mat_mul(mat_mul(mat_translate(t), mat_from_quat(q[0], q[1], q[2], q[3])), mat_scale(s));
}

// ------------------------------------------------------------------------------------------
// a YAML subset: block mappings, block sequences, flow sequences, plain/quoted scalars
// ------------------------------------------------------------------------------------------
struct Node {
    enum Kind { NIL, SCALAR, MAP, SEQ } kind = NIL;
    std::string text;
    std::vector<std::pair<std::string, Node>> map;
    std::vector<Node> seq;

    const Node* get(const char* key) const {
        if (kind != MAP) return nullptr;
        for (const auto& kv : map)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
};

struct Line {
    int indent;
    std::string text;
    int number;
};

std::string trim(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && (s[a] == ' ' || s[a] == '\t' || s[a] == '\r')) ++a;
    while (b > a && (s[b - 1] == ' ' || s[b - 1] == '\t' || s[b - 1] == '\r')) --b;
    return s.substr(a, b - a);
}

struct ParseError {
    std::string what;
};

[[noreturn]] void parse_fail(int line, const char* fmt, ...) {
    char buf[256];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    throw ParseError{ "line " + std::to_string(line) + ": " + buf };
}

Node parse_inline(const std::string& raw, int line);

// splits the inside of a flow collection at top-level commas
std::vector<std::string> split_flow(const std::string& s, int line) {
    std::vector<std::string> out;
    int depth = 0;
    char quote = 0;
    std::string cur;
    for (char ch : s) {
        if (quote) {
            cur += ch;
            if (ch == quote) quote = 0;
            continue;
        }
        if (ch == '"' || ch == '\'') {
            quote = ch;
            cur += ch;
        } else if (ch == '[' || ch == '{') {
            ++depth;
            cur += ch;
        } else if (ch == ']' || ch == '}') {
            --depth;
            cur += ch;
        } else if (ch == ',' && depth == 0) {
            out.push_back(cur);
            cur.clear();
        } else {
            cur += ch;
        }
    }
    if (depth != 0 || quote) parse_fail(line, "unbalanced flow collection");
    if (!trim(cur).empty() || !out.empty()) out.push_back(cur);
    return out;
}

// position of the ": " (or trailing ':') that ends a plain mapping key, or npos
size_t key_colon(const std::string& s) {
    if (s.empty() || s[0] == '[' || s[0] == '{' || s[0] == '"' || s[0] == '\'') return std::string::npos;
    for (size_t i = 0; i < s.size(); ++i) {
        if (s[i] == ':' && (i + 1 == s.size() || s[i + 1] == ' ')) return i;
    }
    return std::string::npos;
}

Node parse_inline(const std::string& raw, int line) {
    const std::string s = trim(raw);
    Node n;
    if (s.empty() || s == "~" || s == "null") return n;
    if (s.front() == '[') {
        if (s.back() != ']') parse_fail(line, "flow sequence does not end on its line");
        n.kind = Node::SEQ;
        for (const std::string& part : split_flow(s.substr(1, s.size() - 2), line)) n.seq.push_back(parse_inline(part, line));
        return n;
    }
    if (s.front() == '{') {
        if (s.back() != '}') parse_fail(line, "flow mapping does not end on its line");
        n.kind = Node::MAP;
        for (const std::string& part : split_flow(s.substr(1, s.size() - 2), line)) {
            const std::string p = trim(part);
            const size_t c = key_colon(p);
            if (c == std::string::npos) parse_fail(line, "flow mapping entry without a key");
            n.map.emplace_back(trim(p.substr(0, c)), parse_inline(p.substr(c + 1), line));
        }
        return n;
    }
    n.kind = Node::SCALAR;
    if ((s.front() == '"' || s.front() == '\'') && s.size() >= 2 && s.back() == s.front()) {
        n.text = s.substr(1, s.size() - 2);
    } else {
        n.text = s;
    }
    return n;
}

Node parse_block(std::vector<Line>& lines, size_t& pos, int indent);

Node parse_value_after_key(std::vector<Line>& lines, size_t& pos, int indent, const std::string& rest, int line) {
    if (!trim(rest).empty()) return parse_inline(rest, line);
    if (pos < lines.size()) {
        const Line& nx = lines[pos];
        const bool is_item = nx.text[0] == '-' && (nx.text.size() == 1 || nx.text[1] == ' ');
        if (nx.indent > indent && !is_item && key_colon(nx.text) == std::string::npos) {  // the value on its own line (`key:\n  {}`)
            ++pos;
            return parse_inline(nx.text, nx.number);
        }
        if (nx.indent > indent || (nx.indent == indent && is_item)) return parse_block(lines, pos, nx.indent);
    }
    return Node();
}

Node parse_block(std::vector<Line>& lines, size_t& pos, int indent) {
    Node n;
    const bool seq = lines[pos].text[0] == '-' && (lines[pos].text.size() == 1 || lines[pos].text[1] == ' ');
    n.kind = seq ? Node::SEQ : Node::MAP;
    while (pos < lines.size() && lines[pos].indent == indent) {
        Line& ln = lines[pos];
        const bool is_item = ln.text[0] == '-' && (ln.text.size() == 1 || ln.text[1] == ' ');
        if (is_item != seq) {
            if (seq) break;  // a sequence that sat at its key's indent ends at the next key
            parse_fail(ln.number, "sequence item inside a mapping");
        }
        if (seq) {
            size_t skip = 1;
            while (skip < ln.text.size() && ln.text[skip] == ' ') ++skip;
            const std::string item = ln.text.substr(skip);
            if (item.empty()) {
                ++pos;
                if (pos < lines.size() && lines[pos].indent > indent) n.seq.push_back(parse_block(lines, pos, lines[pos].indent));
                else n.seq.push_back(Node());
            } else if (key_colon(item) != std::string::npos) {
                ln.indent = indent + (int)skip;  // "- key: v" opens a mapping whose keys sit under the item text
                ln.text = item;
                n.seq.push_back(parse_block(lines, pos, ln.indent));
            } else {
                n.seq.push_back(parse_inline(item, ln.number));
                ++pos;
            }
        } else {
            const size_t c = key_colon(ln.text);
            if (c == std::string::npos) parse_fail(ln.number, "expected `key: value`");
            const std::string key = trim(ln.text.substr(0, c));
            const std::string rest = ln.text.substr(c + 1);
            const int number = ln.number;
            ++pos;
            n.map.emplace_back(key, parse_value_after_key(lines, pos, indent, rest, number));
        }
    }
    if (pos < lines.size() && lines[pos].indent > indent) parse_fail(lines[pos].number, "unexpected indentation");
    return n;
}

Node parse_yaml(const std::string& text) {
    std::vector<Line> lines;
    size_t a = 0;
    int number = 0;
    while (a <= text.size()) {
        size_t b = text.find('\n', a);
        if (b == std::string::npos) b = text.size();
        std::string raw = text.substr(a, b - a);
        a = b + 1;
        ++number;
        int indent = 0;
        while ((size_t)indent < raw.size() && raw[indent] == ' ') ++indent;
        const std::string body = trim(raw);
        if (body.empty() || body[0] == '#' || body == "---" || body == "...") continue;
        if (raw.find('\t') != std::string::npos && raw.find('\t') < (size_t)indent + 1) parse_fail(number, "tab in indentation");
        lines.push_back({ indent, body, number });
    }
    if (lines.empty()) return Node();
    size_t pos = 0;
    Node root = parse_block(lines, pos, lines[0].indent);
    if (pos != lines.size()) parse_fail(lines[pos].number, "content after the end of the document");
    return root;
}

bool read_file(const std::string& path, std::string& out) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? (size_t)n : 0);
    const size_t got = n > 0 ? std::fread(&out[0], 1, (size_t)n, f) : 0;
    std::fclose(f);
    return got == out.size();
}

bool to_float(const Node* n, float& out) {
    if (!n || n->kind != Node::SCALAR) return false;
    const std::string& s = n->text;
    if (s == ".inf" || s == "+.inf" || s == ".Inf") { out = HUGE_VALF; return true; }
    if (s == "-.inf" || s == "-.Inf") { out = -HUGE_VALF; return true; }
    if (s == ".nan" || s == ".NaN") { out = std::strtof("nan", nullptr); return true; }
    char* end = nullptr;
    errno = 0;
    out = std::strtof(s.c_str(), &end);  // correctly rounded decimal -> fp32, as yaml-cpp's stream extraction yields
    return end && *end == 0 && end != s.c_str();
}

bool to_floats(const Node* n, float* out, int count) {
    if (!n || n->kind != Node::SEQ || (int)n->seq.size() < count) return false;
    for (int i = 0; i < count; ++i)
        if (!to_float(&n->seq[i], out[i])) return false;
    return true;
}

bool to_int(const Node* n, long long& out) {
    if (!n || n->kind != Node::SCALAR) return false;
    char* end = nullptr;
    out = std::strtoll(n->text.c_str(), &end, 10);
    return end && *end == 0 && end != n->text.c_str();
}

std::string upper(std::string s) {
    for (char& c : s) c = (char)std::toupper((unsigned char)c);
    return s;
}

// ------------------------------------------------------------------------------------------
// the parsed scene
// ------------------------------------------------------------------------------------------
struct MeshData {
    std::vector<float> positions, normals, texcoords;
    std::vector<uint32_t> indices;
    std::vector<uint32_t> subset_ranges;  // (index_offset, index_count) pairs
    uint32_t flags = 0;
    std::string path, guid;
};

}  // namespace

struct cavept_scene_file {
    std::vector<MeshData> mesh_data;
    std::vector<cavept_mesh> meshes;
    std::vector<cavept_instance> instances;
    std::vector<cavept_material> materials;
    std::vector<long long> instance_entities, material_entities;
    int skipped = 0;
    std::string error;

    void finalize() {
        meshes.clear();
        for (const MeshData& m : mesh_data) {
            cavept_mesh cm;
            cm.positions = m.positions.data();
            cm.normals = m.normals.data();
            cm.indices = m.indices.data();
            cm.vertex_count = (int)(m.positions.size() / 3);
            cm.index_count = (int)m.indices.size();
            meshes.push_back(cm);
        }
    }
};

namespace {

int set_error(char* err, size_t cap, int code, const std::string& what) {
    if (err && cap) {
        std::snprintf(err, cap, "%s", what.c_str());
    }
    return code;
}

// MeshAsset::SerializeBinary read side (mesh_asset.cpp:45-61): u64 subset_count; vector<MeshSubset>
// (u64 n + n x {u32 index_offset, u32 index_count, AABB 2 x vec3}); u32 flags; then the vectors
// indices u32, positions vec3, normals vec3, tangents vec3, texcoords_0 vec2, texcoords_1 vec2,
// joints_0 ivec4, weights_0 vec4, color_0 vec4 -- each u64 count + raw elements.
bool load_mesh_file(const std::string& path, MeshData& out, std::string& why) {
    std::string b;
    if (!read_file(path, b)) {
        why = "cannot read " + path;
        return false;
    }
    size_t o = 0;
    auto need = [&](size_t n) { return o + n >= o && o + n <= b.size(); };
    auto rd_u64 = [&](uint64_t& v) {
        if (!need(8)) return false;
        std::memcpy(&v, b.data() + o, 8);
        o += 8;
        return true;
    };
    auto vec = [&](size_t elem, const char* name, const char** data, uint64_t& count) {
        if (!rd_u64(count) || count > b.size() / elem || !need((size_t)count * elem)) {
            why = path + ": truncated or corrupt `" + name + "` vector";
            return false;
        }
        *data = b.data() + o;
        o += (size_t)count * elem;
        return true;
    };
    uint64_t subset_count = 0, n = 0;
    const char* p = nullptr;
    if (!rd_u64(subset_count)) {
        why = path + ": too short for a .mesh";
        return false;
    }
    if (!vec(32, "subsets", &p, n)) return false;
    if (n != subset_count) {
        why = path + ": subset count mismatch (not a MeshAsset archive?)";
        return false;
    }
    out.subset_ranges.resize((size_t)n * 2);
    for (uint64_t i = 0; i < n; ++i) std::memcpy(&out.subset_ranges[i * 2], p + i * 32, 8);
    if (!need(4)) {
        why = path + ": truncated before flags";
        return false;
    }
    std::memcpy(&out.flags, b.data() + o, 4);
    o += 4;
    if (!vec(4, "indices", &p, n)) return false;
    out.indices.resize((size_t)n);
    if (n) std::memcpy(out.indices.data(), p, (size_t)n * 4);
    if (!vec(12, "positions", &p, n)) return false;
    out.positions.resize((size_t)n * 3);
    if (n) std::memcpy(out.positions.data(), p, (size_t)n * 12);
    if (!vec(12, "normals", &p, n)) return false;
    out.normals.resize((size_t)n * 3);
    if (n) std::memcpy(out.normals.data(), p, (size_t)n * 12);
    if (!vec(12, "tangents", &p, n)) return false;
    if (!vec(8, "texcoords_0", &p, n)) return false;
    out.texcoords.resize((size_t)n * 2);
    if (n) std::memcpy(out.texcoords.data(), p, (size_t)n * 8);
    if (!vec(8, "texcoords_1", &p, n) || !vec(16, "joints_0", &p, n) || !vec(16, "weights_0", &p, n) || !vec(16, "color_0", &p, n)) return false;
    // bytes after color_0 are ignored, as Archive::Read never looks at them
    if (out.indices.empty() || out.indices.size() % 3 != 0) {
        why = path + ": index count is not a positive multiple of 3 (DEV_ASSERT in bvh_accel.cpp:269)";
        return false;
    }
    if (out.positions.empty() || out.normals.size() != out.positions.size()) {
        why = path + ": the path tracer reads positions[i] and normals[i] for every vertex (path_tracer.cpp:33-40); counts differ";
        return false;
    }
    const uint32_t nv = (uint32_t)(out.positions.size() / 3);
    for (uint32_t idx : out.indices) {
        if (idx >= nv) {
            why = path + ": vertex index out of range";
            return false;
        }
    }
    out.path = path;
    return true;
}

bool ends_with(const std::string& s, const char* suffix) {
    const size_t n = std::strlen(suffix);
    return s.size() >= n && s.compare(s.size() - n, n, suffix) == 0;
}

// guid -> asset path for every `*.mesh.meta` below `root`
void scan_meta(const std::string& root, std::map<std::string, std::string>& out, int depth = 0) {
    if (depth > 16) return;
    DIR* d = opendir(root.c_str());
    if (!d) return;
    std::vector<std::string> names;
    while (dirent* e = readdir(d)) names.push_back(e->d_name);
    closedir(d);
    std::sort(names.begin(), names.end());
    for (const std::string& name : names) {
        if (name == "." || name == "..") continue;
        const std::string full = root + "/" + name;
        struct stat st;
        if (stat(full.c_str(), &st) != 0) continue;
        if (S_ISDIR(st.st_mode)) {
            scan_meta(full, out, depth + 1);
        } else if (ends_with(name, ".mesh.meta")) {
            std::string text;
            if (!read_file(full, text)) continue;
            try {
                const Node meta = parse_yaml(text);
                const Node* guid = meta.get("guid");
                if (guid && guid->kind == Node::SCALAR) out.emplace(upper(guid->text), full.substr(0, full.size() - 5));
            } catch (const ParseError&) {
            }
        }
    }
}

struct Entity {
    long long id = 0;
    bool has_transform = false, has_hierarchy = false, has_renderer = false, has_material = false;
    float t[3] = { 0, 0, 0 }, q[4] = { 0, 0, 0, 1 }, s[3] = { 1, 1, 1 };
    long long parent = 0;
    std::string mesh_guid;
    std::vector<long long> materials;
    cavept_material material{};
    int line = 0;
};

}  // namespace

extern "C" {

CAVEPT_API int cavept_mesh_file_open(const char* mesh_path, cavept_scene_file** out, char* err, size_t err_capacity) {
    if (!mesh_path || !out) return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    cavept_scene_file* sf = new cavept_scene_file();
    sf->mesh_data.emplace_back();
    std::string why;
    if (!load_mesh_file(mesh_path, sf->mesh_data[0], why)) {
        delete sf;
        return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, why);
    }
    cavept_instance inst{};
    const Mat4 identity = mat_identity();
    std::memcpy(inst.transform, identity.m, sizeof inst.transform);
    sf->instances.push_back(inst);
    sf->instance_entities.push_back(0);
    // MaterialComponent defaults, cave/engine/scene/material_component.h:12-21
    sf->materials.push_back(cavept_material{ { 1.0f, 1.0f, 1.0f }, 1.0f, 0.0f, 0.0f });
    sf->material_entities.push_back(0);
    sf->finalize();
    *out = sf;
    return CAVEPT_OK;
}

CAVEPT_API int cavept_scene_file_open(const char* scene_path, const char* resource_root, cavept_scene_file** out, char* err, size_t err_capacity) {
    if (!scene_path || !out) return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, "null argument");
    *out = nullptr;
    std::string text;
    if (!read_file(scene_path, text)) return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, std::string("cannot read ") + scene_path);
    Node doc;
    try {
        doc = parse_yaml(text);
    } catch (const ParseError& e) {
        return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, std::string(scene_path) + ": " + e.what);
    }
    const Node* ents = doc.get("entities");
    if (!ents || ents->kind != Node::SEQ) return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, std::string(scene_path) + ": no `entities` list");

    std::vector<Entity> entities;
    std::map<long long, size_t> by_id;
    for (const Node& en : ents->seq) {
        Entity e;
        if (!to_int(en.get("id"), e.id)) return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, std::string(scene_path) + ": entity without an integer id");
        const std::string tag = "entity " + std::to_string(e.id);
        if (const Node* tc = en.get("TransformComponent")) {
            e.has_transform = true;
            // absent keys keep the constructor defaults (transform_component.cpp:8-14)
            if ((tc->get("translation") && !to_floats(tc->get("translation"), e.t, 3)) || (tc->get("rotation") && !to_floats(tc->get("rotation"), e.q, 4)) ||
                (tc->get("scale") && !to_floats(tc->get("scale"), e.s, 3))) {
                return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, tag + ": malformed TransformComponent");
            }
        }
        if (const Node* hc = en.get("HierarchyComponent")) {
            e.has_hierarchy = true;
            if (!to_int(hc->get("parent_id"), e.parent)) return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, tag + ": HierarchyComponent without parent_id");
        }
        if (const Node* mr = en.get("MeshRendererComponent")) {
            e.has_renderer = true;
            const Node* guid = mr->get("mesh_id");
            if (guid && guid->kind == Node::SCALAR) e.mesh_guid = upper(guid->text);
            if (const Node* mats = mr->get("materials")) {
                if (mats->kind == Node::SEQ) {
                    for (const Node& m : mats->seq) {
                        long long id = 0;
                        if (!to_int(&m, id)) return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, tag + ": non-integer material entity");
                        e.materials.push_back(id);
                    }
                }
            }
        }
        if (const Node* mc = en.get("MaterialComponent")) {
            e.has_material = true;
            float bc[4] = { 1, 1, 1, 1 };
            e.material = cavept_material{ { 1.0f, 1.0f, 1.0f }, 1.0f, 0.0f, 0.0f };
            if (mc->get("base_color")) {
                if (!to_floats(mc->get("base_color"), bc, 3)) return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, tag + ": malformed base_color");
                std::memcpy(e.material.base_color, bc, 12);
            }
            if ((mc->get("roughness") && !to_float(mc->get("roughness"), e.material.roughness)) || (mc->get("metallic") && !to_float(mc->get("metallic"), e.material.metallic)) ||
                (mc->get("emissive") && !to_float(mc->get("emissive"), e.material.emissive))) {
                return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, tag + ": malformed MaterialComponent");
            }
        }
        by_id[e.id] = entities.size();
        entities.push_back(e);
    }

    std::string root = resource_root ? resource_root : "";
    if (root.empty()) {
        root = scene_path;
        const size_t slash = root.find_last_of('/');
        root = slash == std::string::npos ? "." : root.substr(0, slash);
    }
    std::map<std::string, std::string> guid_to_path;
    scan_meta(root, guid_to_path);

    cavept_scene_file* sf = new cavept_scene_file();
    std::map<std::string, int> mesh_index;
    std::map<long long, int> material_index;
    for (const Entity& e : entities) {
        if (!e.has_renderer) continue;
        if (!e.has_transform) {  // DEV_VERIFY(transform && mesh) fails -> the entity is skipped (path_tracer.cpp:216)
            ++sf->skipped;
            continue;
        }
        auto path_it = guid_to_path.find(e.mesh_guid);
        if (path_it == guid_to_path.end()) {  // mesh handle does not resolve -> skipped as above
            ++sf->skipped;
            continue;
        }
        auto mi = mesh_index.find(e.mesh_guid);
        if (mi == mesh_index.end()) {
            MeshData md;
            std::string why;
            if (!load_mesh_file(path_it->second, md, why)) {
                delete sf;
                return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, why);
            }
            md.guid = e.mesh_guid;
            mi = mesh_index.emplace(e.mesh_guid, (int)sf->mesh_data.size()).first;
            sf->mesh_data.push_back(std::move(md));
        }
        if (e.materials.empty()) {  // GetMaterialInstances()[0] on an empty vector: undefined in the reference, an error here
            delete sf;
            return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, "entity " + std::to_string(e.id) + ": MeshRendererComponent without materials");
        }
        const long long mat_entity = e.materials[0];
        auto mat_it = material_index.find(mat_entity);
        if (mat_it == material_index.end()) {
            auto ent_it = by_id.find(mat_entity);
            if (ent_it == by_id.end() || !entities[ent_it->second].has_material) {  // DEV_ASSERT(material), path_tracer.cpp:136
                delete sf;
                return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG,
                                 "entity " + std::to_string(e.id) + ": material entity " + std::to_string(mat_entity) + " has no MaterialComponent");
            }
            mat_it = material_index.emplace(mat_entity, (int)sf->materials.size()).first;
            sf->materials.push_back(entities[ent_it->second].material);
            sf->material_entities.push_back(mat_entity);
        }
        // world matrix: UpdateHierarchy for entities with a HierarchyComponent, else the local matrix
        Mat4 world = local_matrix(e.t, e.q, e.s);
        if (e.has_hierarchy) {
            long long parent = e.parent;
            size_t hops = 0;
            while (parent != 0) {
                auto pit = by_id.find(parent);
                if (pit == by_id.end() || !entities[pit->second].has_transform) break;  // DEV_VERIFY(parent_transform) fails -> stop
                const Entity& pe = entities[pit->second];
                world = mat_mul(local_matrix(pe.t, pe.q, pe.s), world);
                parent = pe.has_hierarchy ? pe.parent : 0;
                if (++hops > entities.size()) {
                    delete sf;
                    return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG, "entity " + std::to_string(e.id) + ": HierarchyComponent cycle");
                }
            }
        }
        cavept_instance inst{};
        inst.mesh = mi->second;
        inst.material = mat_it->second;
        std::memcpy(inst.transform, world.m, sizeof inst.transform);
        sf->instances.push_back(inst);
        sf->instance_entities.push_back(e.id);
    }
    if (sf->instances.empty()) {
        const int skipped = sf->skipped;
        delete sf;
        return set_error(err, err_capacity, CAVEPT_ERR_INVALID_ARG,
                         std::string(scene_path) + ": no renderable entity resolves to a .mesh under " + root + " (" + std::to_string(skipped) + " skipped)");
    }
    sf->finalize();
    *out = sf;
    return CAVEPT_OK;
}

CAVEPT_API void cavept_scene_file_close(cavept_scene_file* scene) { delete scene; }

CAVEPT_API int cavept_scene_file_counts(const cavept_scene_file* scene, int* meshes, int* instances, int* materials, int* skipped_renderers) {
    if (!scene) return CAVEPT_ERR_INVALID_ARG;
    if (meshes) *meshes = (int)scene->meshes.size();
    if (instances) *instances = (int)scene->instances.size();
    if (materials) *materials = (int)scene->materials.size();
    if (skipped_renderers) *skipped_renderers = scene->skipped;
    return CAVEPT_OK;
}

CAVEPT_API const cavept_mesh* cavept_scene_file_meshes(const cavept_scene_file* scene) { return scene ? scene->meshes.data() : nullptr; }
CAVEPT_API const cavept_instance* cavept_scene_file_instances(const cavept_scene_file* scene) { return scene ? scene->instances.data() : nullptr; }
CAVEPT_API const cavept_material* cavept_scene_file_materials(const cavept_scene_file* scene) { return scene ? scene->materials.data() : nullptr; }

CAVEPT_API int cavept_upload_scene_file(cavept_ctx* ctx, const cavept_scene_file* scene) {
    if (!ctx || !scene) return CAVEPT_ERR_INVALID_ARG;
    return cavept_upload_scene(ctx, scene->meshes.data(), (int)scene->meshes.size(), scene->instances.data(), (int)scene->instances.size(), scene->materials.data(),
                               (int)scene->materials.size());
}

// MeshAsset::SerializeBinary write side: one subset spanning every index, RENDERABLE flag,
// empty tangent/texcoord/skin/colour vectors.  Lets a procedural mesh be handed to the engine.
CAVEPT_API int cavept_mesh_file_write(const char* path, const cavept_mesh* mesh) {
    if (!path || !mesh || !mesh->positions || !mesh->normals || !mesh->indices || mesh->vertex_count <= 0 || mesh->index_count <= 0 || mesh->index_count % 3 != 0) {
        return CAVEPT_ERR_INVALID_ARG;
    }
    FILE* f = std::fopen(path, "wb");
    if (!f) return CAVEPT_ERR_INVALID_ARG;
    bool ok = true;
    auto put = [&](const void* p, size_t n) { ok = ok && (n == 0 || std::fwrite(p, 1, n, f) == n); };
    auto put_u64 = [&](uint64_t v) { put(&v, 8); };
    float lo[3] = { HUGE_VALF, HUGE_VALF, HUGE_VALF }, hi[3] = { -HUGE_VALF, -HUGE_VALF, -HUGE_VALF };
    for (int i = 0; i < mesh->index_count; ++i) {  // MeshAsset::CreateRenderData's subset bound (mesh_asset.cpp:18-27) before MakeValid
        const float* p = mesh->positions + 3 * (size_t)mesh->indices[i];
        for (int a = 0; a < 3; ++a) {
            lo[a] = p[a] < lo[a] ? p[a] : lo[a];
            hi[a] = p[a] > hi[a] ? p[a] : hi[a];
        }
    }
    put_u64(1);
    put_u64(1);
    const uint32_t range[2] = { 0u, (uint32_t)mesh->index_count };
    put(range, 8);
    put(lo, 12);
    put(hi, 12);
    const uint32_t flags = 1u;  // RENDERABLE = BIT(1), and Bit(x) = (1 << x) >> 1 (cave/engine/math/math.h:10-12)
    put(&flags, 4);
    put_u64((uint64_t)mesh->index_count);
    put(mesh->indices, (size_t)mesh->index_count * 4);
    put_u64((uint64_t)mesh->vertex_count);
    put(mesh->positions, (size_t)mesh->vertex_count * 12);
    put_u64((uint64_t)mesh->vertex_count);
    put(mesh->normals, (size_t)mesh->vertex_count * 12);
    for (int k = 0; k < 6; ++k) put_u64(0);  // tangents, texcoords_0/1, joints_0, weights_0, color_0
    ok = (std::fclose(f) == 0) && ok;
    return ok ? CAVEPT_OK : CAVEPT_ERR_INVALID_ARG;
}

}  // extern "C"
