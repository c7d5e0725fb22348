// Collider ingestion from .glb files — see gltf_colliders.h / include/b2col_gltf.h.
//
// What the reference does (src/gltf_loader.cpp), restated for the collision path only:
//   :112      the file must be a binary glTF container (tinygltf LoadBinaryFromFile)
//   :136-143  scene = `scene` property if present, else scene 0
//   :352-588  per mesh, per primitive WITH a POSITION attribute: AABB = POSITION accessor min / max, cast to
//             float (:576-583); primitives without POSITION are skipped; a primitive without NORMAL (:381-384) or
//             TEXCOORD_0 (:409-412) makes the whole load throw
//   :599      file-root entity named after the file, rotation quat(w,x,y,z) = (1/sqrt2, 1/sqrt2, 0, 0)
//   :603-679  depth-first over the scene's nodes: one entity per node ("anode" if unnamed), TRS from `matrix`
//             (decomposed, :51-77) or from translation / rotation / scale; a mesh with ONE kept primitive puts the
//             collider on the node's entity, with several it creates one child "_mesh<i>" per primitive (:647-669);
//             then the node's children
#include "gltf_colliders.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <utility>

namespace b2gltf {
namespace {

// ---- a small JSON reader (objects keep insertion order; numbers are doubles, like the reference's parser) -----
struct Json {
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Json> items;                           // Array
    std::vector<std::pair<std::string, Json>> fields;  // Object

    const Json* get(const char* key) const
    {
        if (kind != Object) return nullptr;
        for (const auto& f : fields)
            if (f.first == key) return &f.second;
        return nullptr;
    }
    bool is_array() const { return kind == Array; }
    size_t size() const { return items.size(); }
};

class JsonReader {
   public:
    JsonReader(const char* p, size_t n) : begin_(p), p_(p), end_(p + n) {}
    bool Parse(Json& out, std::string& err)
    {
        if (!Value(out, 0)) {
            err = "JSON: " + err_ + " at byte " + std::to_string(static_cast<size_t>(p_ - begin_));
            return false;
        }
        Skip();
        return true;
    }

   private:
    const char* begin_;
    const char* p_;
    const char* end_;
    std::string err_;
    bool Fail(const char* m)
    {
        if (err_.empty()) err_ = m;
        return false;
    }
    void Skip()
    {
        while (p_ < end_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) ++p_;
    }
    bool Lit(const char* s)
    {
        const size_t n = std::strlen(s);
        if (static_cast<size_t>(end_ - p_) < n || std::memcmp(p_, s, n) != 0) return Fail("bad literal");
        p_ += n;
        return true;
    }
    static void Utf8(uint32_t cp, std::string& s)
    {
        if (cp < 0x80) s += static_cast<char>(cp);
        else if (cp < 0x800) {
            s += static_cast<char>(0xC0 | (cp >> 6));
            s += static_cast<char>(0x80 | (cp & 0x3F));
        }
        else if (cp < 0x10000) {
            s += static_cast<char>(0xE0 | (cp >> 12));
            s += static_cast<char>(0x80 | ((cp >> 6) & 0x3F));
            s += static_cast<char>(0x80 | (cp & 0x3F));
        }
        else {
            s += static_cast<char>(0xF0 | (cp >> 18));
            s += static_cast<char>(0x80 | ((cp >> 12) & 0x3F));
            s += static_cast<char>(0x80 | ((cp >> 6) & 0x3F));
            s += static_cast<char>(0x80 | (cp & 0x3F));
        }
    }
    bool Hex4(uint32_t& v)
    {
        if (end_ - p_ < 4) return Fail("short \\u escape");
        v = 0;
        for (int i = 0; i < 4; ++i) {
            const char c = *p_++;
            v <<= 4;
            if (c >= '0' && c <= '9') v |= static_cast<uint32_t>(c - '0');
            else if (c >= 'a' && c <= 'f') v |= static_cast<uint32_t>(c - 'a' + 10);
            else if (c >= 'A' && c <= 'F') v |= static_cast<uint32_t>(c - 'A' + 10);
            else return Fail("bad \\u escape");
        }
        return true;
    }
    bool Str(std::string& s)
    {
        if (p_ >= end_ || *p_ != '"') return Fail("expected string");
        ++p_;
        while (p_ < end_ && *p_ != '"') {
            char c = *p_++;
            if (c != '\\') {
                s += c;
                continue;
            }
            if (p_ >= end_) return Fail("dangling escape");
            c = *p_++;
            switch (c) {
                case '"': s += '"'; break;
                case '\\': s += '\\'; break;
                case '/': s += '/'; break;
                case 'b': s += '\b'; break;
                case 'f': s += '\f'; break;
                case 'n': s += '\n'; break;
                case 'r': s += '\r'; break;
                case 't': s += '\t'; break;
                case 'u': {
                    uint32_t cp;
                    if (!Hex4(cp)) return false;
                    if (cp >= 0xD800 && cp < 0xDC00 && end_ - p_ >= 6 && p_[0] == '\\' && p_[1] == 'u') {
                        p_ += 2;
                        uint32_t lo;
                        if (!Hex4(lo)) return false;
                        cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                    }
                    Utf8(cp, s);
                    break;
                }
                default: return Fail("bad escape");
            }
        }
        if (p_ >= end_) return Fail("unterminated string");
        ++p_;
        return true;
    }
    bool Value(Json& v, int depth);
};

bool JsonReader::Value(Json& v, int depth)
{
    if (depth > 256) return Fail("nesting too deep");
    Skip();
    if (p_ >= end_) return Fail("unexpected end");
    const char c = *p_;
    if (c == '{') {
        ++p_;
        v.kind = Json::Object;
        Skip();
        if (p_ < end_ && *p_ == '}') {
            ++p_;
            return true;
        }
        for (;;) {
            Skip();
            std::string key;
            if (!Str(key)) return false;
            Skip();
            if (p_ >= end_ || *p_ != ':') return Fail("expected ':'");
            ++p_;
            v.fields.emplace_back(std::move(key), Json());
            if (!Value(v.fields.back().second, depth + 1)) return false;
            Skip();
            if (p_ < end_ && *p_ == ',') {
                ++p_;
                continue;
            }
            if (p_ < end_ && *p_ == '}') {
                ++p_;
                return true;
            }
            return Fail("expected ',' or '}'");
        }
    }
    if (c == '[') {
        ++p_;
        v.kind = Json::Array;
        Skip();
        if (p_ < end_ && *p_ == ']') {
            ++p_;
            return true;
        }
        for (;;) {
            v.items.emplace_back();
            if (!Value(v.items.back(), depth + 1)) return false;
            Skip();
            if (p_ < end_ && *p_ == ',') {
                ++p_;
                continue;
            }
            if (p_ < end_ && *p_ == ']') {
                ++p_;
                return true;
            }
            return Fail("expected ',' or ']'");
        }
    }
    if (c == '"') {
        v.kind = Json::String;
        return Str(v.str);
    }
    if (c == 't') {
        v.kind = Json::Bool;
        v.b = true;
        return Lit("true");
    }
    if (c == 'f') {
        v.kind = Json::Bool;
        v.b = false;
        return Lit("false");
    }
    if (c == 'n') return Lit("null");
    if (c == '-' || (c >= '0' && c <= '9')) {
        // the number's text, then strtod: correctly rounded to double, as the reference's JSON library does
        const char* q = p_;
        while (q < end_ && (*q == '-' || *q == '+' || *q == '.' || *q == 'e' || *q == 'E' || (*q >= '0' && *q <= '9'))) ++q;
        const std::string text(p_, q);
        char* stop = nullptr;
        v.num = std::strtod(text.c_str(), &stop);
        if (stop == text.c_str() || *stop != '\0') return Fail("bad number");
        v.kind = Json::Number;
        p_ = q;
        return true;
    }
    return Fail("unexpected character");
}

// ---- helpers over the glTF document --------------------------------------------------------------------------
bool Index(const Json* v, size_t limit, size_t& out)
{
    if (!v || v->kind != Json::Number || v->num < 0.0 || v->num != std::floor(v->num) || v->num >= static_cast<double>(limit)) return false;
    out = static_cast<size_t>(v->num);
    return true;
}

bool Floats(const Json* v, size_t n, float* out)
{
    if (!v || !v->is_array() || v->size() != n) return false;
    for (size_t i = 0; i < n; ++i) {
        if (v->items[i].kind != Json::Number) return false;
        out[i] = static_cast<float>(v->items[i].num);  // the reference casts each double to float (:577-582, :626-639)
    }
    return true;
}

// glm::quat_cast(mat3) (reference vendor/glm/glm/gtc/quaternion.inl:81-123); m is column-major 4x4, m[c*4+r]
void QuatCast(const float* m, float* q_xyzw)
{
    const float m00 = m[0], m11 = m[5], m22 = m[10];
    const float fx = m00 - m11 - m22, fy = m11 - m00 - m22, fz = m22 - m00 - m11, fw = m00 + m11 + m22;
    int biggest = 0;
    float four = fw;
    if (fx > four) { four = fx; biggest = 1; }
    if (fy > four) { four = fy; biggest = 2; }
    if (fz > four) { four = fz; biggest = 3; }
    const float val = std::sqrt(four + 1.0f) * 0.5f;
    const float mult = 0.25f / val;
    const float m01 = m[1], m02 = m[2], m10 = m[4], m12 = m[6], m20 = m[8], m21 = m[9];
    float w, x, y, z;
    switch (biggest) {
        case 0: w = val; x = (m12 - m21) * mult; y = (m20 - m02) * mult; z = (m01 - m10) * mult; break;
        case 1: w = (m12 - m21) * mult; x = val; y = (m01 + m10) * mult; z = (m20 + m02) * mult; break;
        case 2: w = (m20 - m02) * mult; x = (m01 + m10) * mult; y = val; z = (m12 + m21) * mult; break;
        default: w = (m01 - m10) * mult; x = (m20 + m02) * mult; y = (m12 + m21) * mult; z = val; break;
    }
    q_xyzw[0] = x; q_xyzw[1] = y; q_xyzw[2] = z; q_xyzw[3] = w;
}

// DecomposeTransform (reference src/gltf_loader.cpp:51-77)
void Decompose(float* m, float* pos, float* rot_xyzw, float* scl)
{
    pos[0] = m[12]; pos[1] = m[13]; pos[2] = m[14];
    m[12] = m[13] = m[14] = 0.0f;
    for (int c = 0; c < 3; ++c) scl[c] = sqrtf(m[4 * c] * m[4 * c] + m[4 * c + 1] * m[4 * c + 1] + m[4 * c + 2] * m[4 * c + 2]);
    for (int row = 0; row < 3; ++row) {
        m[row] /= scl[0];
        m[4 + row] /= scl[1];
        m[8 + row] /= scl[2];
    }
    QuatCast(m, rot_xyzw);
}

struct Prim {
    float aabb[6];
};

struct Loader {
    const Json& doc;
    bool strict;
    Scene& out;
    std::string& error;
    std::vector<std::vector<Prim>> mesh_prims;  // kept primitives per mesh (reference `primitive_arrays`, :349)
    const Json* nodes = nullptr;
    size_t n_nodes = 0;
    size_t visited = 0;

    Status Fail(Status st, const std::string& m)
    {
        error = m;
        return st;
    }

    Status Meshes()
    {
        const Json* meshes = doc.get("meshes");
        const Json* accessors = doc.get("accessors");
        const size_t n_acc = accessors && accessors->is_array() ? accessors->size() : 0;
        if (!meshes) return OK;
        if (!meshes->is_array()) return Fail(ERR_FORMAT, "`meshes` is not an array");
        mesh_prims.resize(meshes->size());
        for (size_t mi = 0; mi < meshes->size(); ++mi) {
            const Json& mesh = meshes->items[mi];
            const Json* name = mesh.get("name");
            const std::string mesh_name = name && name->kind == Json::String ? name->str : std::string();
            const Json* prims = mesh.get("primitives");
            if (!prims || !prims->is_array()) return Fail(ERR_FORMAT, "mesh " + std::to_string(mi) + " has no primitives array");
            for (const Json& prim : prims->items) {
                const Json* attrs = prim.get("attributes");
                const Json* position = attrs ? attrs->get("POSITION") : nullptr;
                if (!position) continue;  // "skip primitive's rendering" (:586-589)
                size_t acc;
                if (!Index(position, n_acc, acc)) return Fail(ERR_FORMAT, "POSITION accessor index out of range in mesh " + mesh_name);
                if (strict) {  // the conditions under which the reference throws
                    if (!attrs->get("NORMAL")) return Fail(ERR_UNSUPPORTED, "No normals found in primitive from " + mesh_name);
                    if (!attrs->get("TEXCOORD_0")) return Fail(ERR_UNSUPPORTED, "No TEXCOORD_0 found in primitive from " + mesh_name);
                    size_t idx;
                    if (!Index(prim.get("indices"), n_acc, idx)) return Fail(ERR_UNSUPPORTED, "primitive without a valid index accessor in " + mesh_name);
                }
                Prim p;
                const Json& a = accessors->items[acc];
                if (!Floats(a.get("min"), 3, p.aabb) || !Floats(a.get("max"), 3, p.aabb + 3))
                    return Fail(ERR_FORMAT, "POSITION accessor without 3-component min/max in mesh " + mesh_name);  // the reference's .at(0..2) would throw
                mesh_prims[mi].push_back(p);
            }
        }
        return OK;
    }

    uint32_t NewEntity(const std::string& name, uint32_t parent)
    {
        Entity e{};
        e.name = name;
        e.parent = parent;
        e.rot[3] = 1.0f;
        e.scl[0] = e.scl[1] = e.scl[2] = 1.0f;
        out.entities.push_back(e);
        return out.first_entity + static_cast<uint32_t>(out.entities.size()) - 1u;
    }
    Entity& At(uint32_t id) { return out.entities[id - out.first_entity]; }

    Status Node(uint32_t parent_entity, size_t ni, int depth)
    {
        if (depth > 512 || ++visited > 16u * n_nodes + 16u) return Fail(ERR_FORMAT, "node hierarchy is cyclic or too deep");
        const Json& node = nodes->items[ni];
        const Json* name = node.get("name");
        const uint32_t id = NewEntity(name && name->kind == Json::String && !name->str.empty() ? name->str : std::string("anode"), parent_entity);
        {
            Entity& e = At(id);
            float m[16];
            if (Floats(node.get("matrix"), 16, m)) {
                Decompose(m, e.pos, e.rot, e.scl);
            }
            else {
                float v[4];
                if (Floats(node.get("translation"), 3, v)) std::memcpy(e.pos, v, 12);
                if (Floats(node.get("rotation"), 4, v)) std::memcpy(e.rot, v, 16);
                if (Floats(node.get("scale"), 3, v)) std::memcpy(e.scl, v, 12);
            }
        }
        if (const Json* mesh = node.get("mesh")) {
            size_t mi;
            if (!Index(mesh, mesh_prims.size(), mi)) return Fail(ERR_FORMAT, "node refers to a mesh outside the array");
            const std::vector<Prim>& prims = mesh_prims[mi];
            if (prims.size() == 1) {
                Entity& e = At(id);
                e.has_collider = true;
                std::memcpy(e.aabb, prims[0].aabb, 24);
            }
            else {
                for (size_t i = 0; i < prims.size(); ++i) {
                    const uint32_t child = NewEntity("_mesh" + std::to_string(i), id);
                    Entity& c = At(child);
                    c.has_collider = true;
                    std::memcpy(c.aabb, prims[i].aabb, 24);
                }
            }
        }
        if (const Json* children = node.get("children")) {
            if (!children->is_array()) return Fail(ERR_FORMAT, "`children` is not an array");
            for (const Json& c : children->items) {
                size_t ci;
                if (!Index(&c, n_nodes, ci)) return Fail(ERR_FORMAT, "child node index out of range");
                const Status st = Node(id, ci, depth + 1);
                if (st != OK) return st;
            }
        }
        return OK;
    }
};

uint32_t ReadU32(const unsigned char* p) { return static_cast<uint32_t>(p[0]) | (static_cast<uint32_t>(p[1]) << 8) | (static_cast<uint32_t>(p[2]) << 16) | (static_cast<uint32_t>(p[3]) << 24); }

}  // namespace

Status LoadMemory(const void* data, size_t size, const std::string& stem, uint32_t first_entity, uint32_t root_parent, bool strict, Scene& out,
                  std::string& error)
{
    out.entities.clear();
    out.first_entity = first_entity ? first_entity : 1u;
    const unsigned char* bytes = static_cast<const unsigned char*>(data);
    // GLB container: 12-byte header (magic "glTF", version, total length), then chunks (length, type, payload);
    // the first chunk is the JSON document
    if (!bytes || size < 20 || std::memcmp(bytes, "glTF", 4) != 0) {
        error = "not a binary glTF container (the reference only loads .glb, src/gltf_loader.cpp:97,112)";
        return ERR_FORMAT;
    }
    if (ReadU32(bytes + 4) != 2u) {
        error = "unsupported glTF container version";
        return ERR_FORMAT;
    }
    const uint32_t json_len = ReadU32(bytes + 12);
    if (ReadU32(bytes + 16) != 0x4E4F534Au || 20ull + json_len > size) {
        error = "first chunk is not a complete JSON chunk";
        return ERR_FORMAT;
    }
    Json doc;
    {
        JsonReader reader(reinterpret_cast<const char*>(bytes + 20), json_len);
        if (!reader.Parse(doc, error)) return ERR_FORMAT;
    }
    if (doc.kind != Json::Object) {
        error = "glTF document is not a JSON object";
        return ERR_FORMAT;
    }
    const Json* scenes = doc.get("scenes");
    if (!scenes || !scenes->is_array() || scenes->size() < 1) {
        error = "glTF file has no scenes";  // the reference throws here (:136-138)
        return ERR_FORMAT;
    }
    size_t scene_index = 0;
    if (const Json* def = doc.get("scene")) {
        if (!Index(def, scenes->size(), scene_index)) {
            error = "default scene index out of range";
            return ERR_FORMAT;
        }
    }
    Loader L{doc, strict, out, error, {}, nullptr, 0, 0};
    L.nodes = doc.get("nodes");
    L.n_nodes = L.nodes && L.nodes->is_array() ? L.nodes->size() : 0;
    Status st = L.Meshes();
    if (st != OK) return st;
    // file-root entity: Y-up -> Z-up, glm::quat{w = 1/sqrt2, x = 1/sqrt2, y = 0, z = 0} (:599)
    const uint32_t root = L.NewEntity(stem, root_parent);
    {
        Entity& r = L.At(root);
        const float one_over_root_two = 0.707106781186547524400844362104849039f;  // glm::one_over_root_two<float>()
        r.rot[0] = one_over_root_two;
        r.rot[3] = one_over_root_two;
    }
    if (const Json* roots = scenes->items[scene_index].get("nodes")) {
        if (!roots->is_array()) {
            error = "scene `nodes` is not an array";
            return ERR_FORMAT;
        }
        for (const Json& n : roots->items) {
            size_t ni;
            if (!Index(&n, L.n_nodes, ni)) {
                error = "scene node index out of range";
                return ERR_FORMAT;
            }
            st = L.Node(root, ni, 0);
            if (st != OK) return st;
        }
    }
    return OK;
}

Status LoadFile(const std::string& path, uint32_t first_entity, uint32_t root_parent, bool strict, Scene& out, std::string& error)
{
    std::FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) {
        error = "cannot open " + path;
        return ERR_IO;
    }
    std::vector<unsigned char> buf;
    unsigned char chunk[65536];
    size_t got;
    while ((got = std::fread(chunk, 1, sizeof chunk, f)) > 0) buf.insert(buf.end(), chunk, chunk + got);
    std::fclose(f);
    // std::filesystem::path(path).stem() (:595-596): the file name up to its last '.'
    std::string stem = path;
    const size_t slash = stem.find_last_of("/\\");
    if (slash != std::string::npos) stem = stem.substr(slash + 1);
    const size_t dot = stem.find_last_of('.');
    if (dot != std::string::npos && dot != 0) stem = stem.substr(0, dot);
    return LoadMemory(buf.data(), buf.size(), stem, first_entity, root_parent, strict, out, error);
}

}  // namespace b2gltf
