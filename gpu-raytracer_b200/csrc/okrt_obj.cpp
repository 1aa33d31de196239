/*
 * okrt_obj.cpp — Wavefront OBJ / MTL reader (row f1 of SURVEY.md §8): what the reference obtains from
 * Importer::loadObjects (source/Graphics/Importer.cpp:77-120) and aiMeshToMeshData (:15-56), i.e. from assimp's
 * OBJ importer run with  aiProcess_Triangulate | aiProcess_ConvertToLeftHanded | aiProcess_CalcTangentSpace .
 *
 * assimp itself is a prebuilt Windows library in the reference tree (deps/lib, no source): nothing of it can be
 * compiled or run here, and no reference test pins its output — parity at the import boundary is UNPINNED.  This
 * file restates assimp 5.x's published behaviour for the path the reference takes:
 *   ObjFileParser      o / g start a new object (groups are mapped onto objects), usemtl starts a new mesh when
 *                      the current one already has faces of another material, faces accept v, v/vt, v//vn, v/vt/vn
 *                      and negative (relative) indices; the mesh is named after the object or the material that
 *                      opened it; meshes without faces are dropped
 *   createVertexArray  every face corner becomes its own vertex (no sharing)
 *   MakeLeftHanded     z -> -z for positions and normals;  FlipUVs  v -> 1 - v;  FlipWindingOrder  reverses
 *                      the corners of every face
 *   Triangulate        triangles pass, quads are split from the (at most one) concave corner, larger polygons are
 *                      fanned from corner 0 (assimp ear-clips those: identical for convex polygons only)
 *   CalcTangentSpace   per-face tangent/bitangent from the UV gradients, Gram-Schmidt against the vertex normal,
 *                      then averaged over vertices at the same position with matching normal (>= 0.9999) and
 *                      tangent frame (within 45 degrees)
 *   materials          map_Kd -> DIFFUSE -> albedo, map_Pr -> DIFFUSE_ROUGHNESS -> roughness, map_Pm -> METALNESS ->
 *                      metallic, map_Ks -> SPECULAR -> specular, norm / map_Kn -> NORMALS -> normal map, else
 *                      disp / map_disp -> DISPLACEMENT -> normal map; map_bump is HEIGHT, which Importer.cpp:99-114
 *                      never asks for
 * followed by Importer.cpp:15-56: position * scale, missing UV / tangents -> 0, bounding box of the scaled points.
 * All arithmetic is float32 (ai_real), compiled with -ffp-contract=off.
 */
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <map>
#include <string>
#include <vector>

#include "okrt.h"
#include "okrt_internal.h"

namespace {

struct V3f { float x, y, z; };
inline V3f operator+(V3f a, V3f b) { return V3f{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3f operator-(V3f a, V3f b) { return V3f{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3f operator*(V3f a, float s) { return V3f{a.x * s, a.y * s, a.z * s}; }
inline float dot3(V3f a, V3f b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3f cross3(V3f a, V3f b) { return V3f{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline float length3(V3f a) { return std::sqrt(dot3(a, a)); }
inline V3f normalizeSafe(V3f a) { const float l = length3(a); return l > 0.f ? a * (1.0f / l) : a; }   /* aiVector3t::NormalizeSafe */
inline V3f normalize3(V3f a) { return a * (1.0f / length3(a)); }                                      /* aiVector3t::Normalize */
inline bool special(float f) { return std::isnan(f) || std::isinf(f); }

struct Corner { int v, vt, vn; };                       /* 0-based, -1 = absent */
struct Face { std::vector<Corner> corners; };
struct ObjMesh { std::string name; int material = -1; std::vector<Face> faces; bool hasNormals = false, hasUvs = false; };
struct ObjObject { std::string name; std::vector<int> meshes; };
struct ObjMaterial { std::string name; std::string tex[5]; std::string texDisplacement; };   /* slots: okrt_obj_texture_slot */

struct LoadedMesh {
    std::string name;
    std::vector<okrt_triangle> pos;
    std::vector<okrt_triangle_info> info;
    okrt_aabb box;
    std::string tex[5];
};

const char* skipSpace(const char* p, const char* e) { while (p < e && (*p == ' ' || *p == '\t' || *p == '\r')) p++; return p; }
const char* tokenEnd(const char* p, const char* e) { while (p < e && *p != ' ' && *p != '\t' && *p != '\r' && *p != '\n') p++; return p; }
std::string restOfLine(const char* p, const char* e)
{
    p = skipSpace(p, e);
    const char* q = p;
    while (q < e && *q != '\n') q++;
    while (q > p && (q[-1] == ' ' || q[-1] == '\t' || q[-1] == '\r')) q--;
    return std::string(p, q);
}
float parseFloat(const char*& p, const char* e)
{
    p = skipSpace(p, e);
    const char* q = tokenEnd(p, e);
    const std::string tok(p, q);
    p = q;
    return tok.empty() ? 0.f : std::strtof(tok.c_str(), nullptr);       /* correctly rounded decimal -> float32, as assimp's fast_atoreal_move is meant to be */
}

struct Parser {
    std::vector<V3f> positions, normals;
    std::vector<V3f> uvs;
    std::vector<ObjMesh> meshes;
    std::vector<ObjObject> objects;
    std::vector<ObjMaterial> materials;
    std::map<std::string, int> materialIndex;
    int currentObject = -1, currentMesh = -1, currentMaterial = -1;
    std::string activeGroup;
    std::vector<std::string> mtlLibs;

    void createMesh(const std::string& name)
    {
        ObjMesh m; m.name = name; m.material = currentMaterial;
        meshes.push_back(m);
        currentMesh = (int)meshes.size() - 1;
        if (currentObject < 0) { ObjObject o; o.name = "defaultobject"; objects.push_back(o); currentObject = 0; }
        objects[currentObject].meshes.push_back(currentMesh);
    }
    void createObject(const std::string& name)
    {
        ObjObject o; o.name = name;
        objects.push_back(o);
        currentObject = (int)objects.size() - 1;
        createMesh(name);
    }
    bool needsNewMesh(int material) const
    {
        if (currentMesh < 0) return true;
        const ObjMesh& m = meshes[currentMesh];
        return m.material != -1 && m.material != material && !m.faces.empty();      /* ObjFileParser::needsNewMesh */
    }

    static int resolveIndex(long idx, size_t count)      /* 1-based, negative = relative to the end */
    {
        if (idx > 0) return (int)idx - 1;
        if (idx < 0) return (int)((long)count + idx);
        return -1;
    }

    void parseFace(const char* p, const char* e)
    {
        Face f;
        bool anyVt = false, anyVn = false;
        for (;;) {
            p = skipSpace(p, e);
            if (p >= e || *p == '\n') break;
            const char* q = tokenEnd(p, e);
            long idx[3] = {0, 0, 0};
            int which = 0;
            const char* c = p;
            while (c < q && which < 3) {
                if (*c == '/') { which++; c++; continue; }
                char* endp = nullptr;
                idx[which] = std::strtol(c, &endp, 10);
                c = endp > c ? endp : c + 1;
            }
            Corner cr;
            cr.v = resolveIndex(idx[0], positions.size());
            cr.vt = resolveIndex(idx[1], uvs.size());
            cr.vn = resolveIndex(idx[2], normals.size());
            if (cr.v < 0 || cr.v >= (int)positions.size()) { p = q; continue; }
            if (cr.vt >= (int)uvs.size()) cr.vt = -1;
            if (cr.vn >= (int)normals.size()) cr.vn = -1;
            anyVt = anyVt || cr.vt >= 0; anyVn = anyVn || cr.vn >= 0;
            f.corners.push_back(cr);
            p = q;
        }
        if (f.corners.empty()) return;
        if (currentMesh < 0) { if (currentObject < 0) createObject("defaultobject"); else createMesh("defaultobject"); }
        ObjMesh& m = meshes[currentMesh];
        if (m.material < 0) m.material = currentMaterial;
        m.hasNormals = m.hasNormals || anyVn;
        m.hasUvs = m.hasUvs || anyVt;
        m.faces.push_back(std::move(f));
    }

    void parse(const char* p, const char* e)
    {
        while (p < e) {
            const char* lineEnd = p;
            while (lineEnd < e && *lineEnd != '\n') lineEnd++;
            const char* s = skipSpace(p, lineEnd);
            const char* t = tokenEnd(s, lineEnd);
            const std::string key(s, t);
            if (key == "v") { V3f v; v.x = parseFloat(t, lineEnd); v.y = parseFloat(t, lineEnd); v.z = parseFloat(t, lineEnd); positions.push_back(v); }
            else if (key == "vn") { V3f v; v.x = parseFloat(t, lineEnd); v.y = parseFloat(t, lineEnd); v.z = parseFloat(t, lineEnd); normals.push_back(v); }
            else if (key == "vt") { V3f v; v.x = parseFloat(t, lineEnd); v.y = parseFloat(t, lineEnd); v.z = parseFloat(t, lineEnd); uvs.push_back(v); }
            else if (key == "f") parseFace(t, lineEnd);
            else if (key == "o") { const std::string name = restOfLine(t, lineEnd); if (!name.empty()) createObject(name); }
            else if (key == "g") { const std::string name = restOfLine(t, lineEnd); if (name != activeGroup) { createObject(name); activeGroup = name; } }
            else if (key == "usemtl") {
                const std::string name = restOfLine(t, lineEnd);
                auto it = materialIndex.find(name);
                int mat;
                if (it == materialIndex.end()) {          /* unknown material: assimp creates a default one under that name */
                    ObjMaterial m; m.name = name; materials.push_back(m);
                    mat = (int)materials.size() - 1; materialIndex[name] = mat;
                } else mat = it->second;
                if (needsNewMesh(mat)) { currentMaterial = mat; createMesh(name); }
                currentMaterial = mat;
                if (currentMesh >= 0) meshes[currentMesh].material = mat;
            }
            else if (key == "mtllib") mtlLibs.push_back(restOfLine(t, lineEnd));
            p = lineEnd < e ? lineEnd + 1 : e;
        }
    }

    static std::string texturePath(const char* p, const char* e)
    {
        /* options (-bm 1.0, -o u v w, -s u v w, -clamp on, ...) precede the file name: keep the last token run that
         * is not an option or an option value */
        std::vector<std::string> toks;
        while (true) { p = skipSpace(p, e); if (p >= e) break; const char* q = tokenEnd(p, e); toks.emplace_back(p, q); p = q; }
        size_t i = 0;
        while (i < toks.size() && toks[i].size() > 1 && toks[i][0] == '-') {
            const std::string& o = toks[i];
            size_t args = 1;
            if (o == "-o" || o == "-s" || o == "-t") args = 3;
            else if (o == "-mm") args = 2;
            i += 1 + args;
        }
        std::string out;
        for (; i < toks.size(); i++) { if (!out.empty()) out += " "; out += toks[i]; }
        std::replace(out.begin(), out.end(), '\\', '/');
        return out;
    }

    void parseMtl(const char* p, const char* e)
    {
        int cur = -1;
        while (p < e) {
            const char* lineEnd = p;
            while (lineEnd < e && *lineEnd != '\n') lineEnd++;
            const char* s = skipSpace(p, lineEnd);
            const char* t = tokenEnd(s, lineEnd);
            std::string key(s, t);
            std::transform(key.begin(), key.end(), key.begin(), [](unsigned char c) { return (char)std::tolower(c); });
            if (key == "newmtl") {
                const std::string name = restOfLine(t, lineEnd);
                auto it = materialIndex.find(name);
                if (it == materialIndex.end()) { ObjMaterial m; m.name = name; materials.push_back(m); cur = (int)materials.size() - 1; materialIndex[name] = cur; }
                else cur = it->second;
            } else if (cur >= 0) {
                const char* te = lineEnd;
                while (te > t && (te[-1] == '\r' || te[-1] == ' ' || te[-1] == '\t')) te--;
                ObjMaterial& m = materials[cur];
                if (key == "map_kd") m.tex[OKRT_OBJ_TEX_ALBEDO] = texturePath(t, te);
                else if (key == "map_pr") m.tex[OKRT_OBJ_TEX_ROUGHNESS] = texturePath(t, te);
                else if (key == "map_pm") m.tex[OKRT_OBJ_TEX_METALLIC] = texturePath(t, te);
                else if (key == "map_ks") m.tex[OKRT_OBJ_TEX_SPECULAR] = texturePath(t, te);
                else if (key == "norm" || key == "map_kn") m.tex[OKRT_OBJ_TEX_NORMAL] = texturePath(t, te);
                else if (key == "disp" || key == "map_disp") m.texDisplacement = texturePath(t, te);
            }
            p = lineEnd < e ? lineEnd + 1 : e;
        }
    }
};

/* TriangulateProcess for one (already de-indexed, left-handed) face: corner indices into the mesh's vertex array */
void triangulate(const std::vector<V3f>& verts, const std::vector<uint32_t>& idx, std::vector<uint32_t>& out)
{
    const size_t n = idx.size();
    if (n < 3) return;                                   /* points and lines are not geometry the reference can use */
    if (n == 3) { out.insert(out.end(), idx.begin(), idx.end()); return; }
    if (n == 4) {
        /* quads have at most one concave corner: start the two-triangle fan there (TriangulateProcess.cpp) */
        uint32_t start = 0;
        for (uint32_t i = 0; i < 4; i++) {
            const V3f v0 = verts[idx[(i + 3) % 4]], v1 = verts[idx[(i + 2) % 4]], v2 = verts[idx[(i + 1) % 4]], v = verts[idx[i]];
            const V3f left = normalize3(v0 - v), diag = normalize3(v1 - v), right = normalize3(v2 - v);
            const float angle = std::acos(dot3(left, diag)) + std::acos(dot3(right, diag));
            if (angle > 3.14159265358979323846f) { start = i; break; }
        }
        const uint32_t t[4] = {idx[0], idx[1], idx[2], idx[3]};
        out.push_back(t[start]); out.push_back(t[(start + 1) % 4]); out.push_back(t[(start + 2) % 4]);
        out.push_back(t[start]); out.push_back(t[(start + 2) % 4]); out.push_back(t[(start + 3) % 4]);
        return;
    }
    for (size_t k = 1; k + 1 < n; k++) { out.push_back(idx[0]); out.push_back(idx[k]); out.push_back(idx[k + 1]); }
}

void calcTangents(const std::vector<V3f>& pos, const std::vector<V3f>& nor, const std::vector<V3f>& uv, const std::vector<uint32_t>& tris,
                  std::vector<V3f>& tang, std::vector<V3f>& bitang)
{
    const size_t nv = pos.size();
    const float qnan = std::numeric_limits<float>::quiet_NaN();
    tang.assign(nv, V3f{qnan, qnan, qnan}); bitang.assign(nv, V3f{qnan, qnan, qnan});
    for (size_t f = 0; f + 2 < tris.size(); f += 3) {
        const uint32_t p0 = tris[f], p1 = tris[f + 1], p2 = tris[f + 2];
        const V3f v = pos[p1] - pos[p0], w = pos[p2] - pos[p0];
        float sx = uv[p1].x - uv[p0].x, sy = uv[p1].y - uv[p0].y;
        float tx = uv[p2].x - uv[p0].x, ty = uv[p2].y - uv[p0].y;
        const float dirCorrection = (tx * sy - ty * sx) < 0.0f ? -1.0f : 1.0f;
        if (sx * ty == sy * tx) { sx = 0.0f; sy = 1.0f; tx = 1.0f; ty = 0.0f; }     /* degenerate UVs: default frame */
        V3f tangent, bitangent;
        tangent.x = (w.x * sy - v.x * ty) * dirCorrection;
        tangent.y = (w.y * sy - v.y * ty) * dirCorrection;
        tangent.z = (w.z * sy - v.z * ty) * dirCorrection;
        bitangent.x = (-w.x * sx + v.x * tx) * dirCorrection;
        bitangent.y = (-w.y * sx + v.y * tx) * dirCorrection;
        bitangent.z = (-w.z * sx + v.z * tx) * dirCorrection;
        for (int b = 0; b < 3; b++) {
            const uint32_t p = tris[f + b];
            V3f lt = tangent - nor[p] * dot3(tangent, nor[p]);
            V3f lb = bitangent - nor[p] * dot3(bitangent, nor[p]) - lt * dot3(bitangent, lt);
            lt = normalizeSafe(lt); lb = normalizeSafe(lb);
            const bool badT = special(lt.x) || special(lt.y) || special(lt.z);
            const bool badB = special(lb.x) || special(lb.y) || special(lb.z);
            if (badT != badB) {
                if (badT) lt = normalizeSafe(cross3(nor[p], lb));
                else lb = normalizeSafe(cross3(lt, nor[p]));
            }
            tang[p] = lt; bitang[p] = lb;
        }
    }
    /* second pass: average the frames of vertices at the same position whose normals and frames agree */
    V3f lo{FLT_MAX, FLT_MAX, FLT_MAX}, hi{-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (const V3f& p : pos) { lo.x = std::min(lo.x, p.x); lo.y = std::min(lo.y, p.y); lo.z = std::min(lo.z, p.z); hi.x = std::max(hi.x, p.x); hi.y = std::max(hi.y, p.y); hi.z = std::max(hi.z, p.z); }
    const float posEpsilon = nv ? length3(hi - lo) * 1e-4f : 0.f;
    const float angleEpsilon = 0.9999f, fLimit = std::cos(45.0f * 3.14159265358979323846f / 180.0f);
    std::vector<uint32_t> order(nv);
    for (uint32_t i = 0; i < nv; i++) order[i] = i;
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return pos[a].x != pos[b].x ? pos[a].x < pos[b].x : a < b; });
    std::vector<uint32_t> rank(nv);
    for (uint32_t i = 0; i < nv; i++) rank[order[i]] = i;
    std::vector<uint8_t> done(nv, 0);
    std::vector<uint32_t> close;
    for (uint32_t a = 0; a < nv; a++) {
        if (done[a]) continue;
        const V3f oPos = pos[a], oNor = nor[a], oT = tang[a], oB = bitang[a];
        close.clear();
        /* candidates: vertices within posEpsilon (sweep along x in the sorted order), visited in ascending index */
        std::vector<uint32_t> found;
        for (uint32_t k = rank[a];; k--) { const uint32_t i = order[k]; if (oPos.x - pos[i].x > posEpsilon) break; if (length3(pos[i] - oPos) < posEpsilon || i == a) found.push_back(i); if (k == 0) break; }
        for (uint32_t k = rank[a] + 1; k < nv; k++) { const uint32_t i = order[k]; if (pos[i].x - oPos.x > posEpsilon) break; if (length3(pos[i] - oPos) < posEpsilon) found.push_back(i); }
        std::sort(found.begin(), found.end());
        for (uint32_t i : found) {
            if (done[i]) continue;
            if (i != a) {
                if (dot3(nor[i], oNor) < angleEpsilon) continue;
                if (dot3(tang[i], oT) < fLimit) continue;
                if (dot3(bitang[i], oB) < fLimit) continue;
            }
            close.push_back(i);
            done[i] = 1;
        }
        V3f st{0, 0, 0}, sb{0, 0, 0};
        for (uint32_t i : close) { st = st + tang[i]; sb = sb + bitang[i]; }
        st = normalize3(st); sb = normalize3(sb);
        for (uint32_t i : close) { tang[i] = st; bitang[i] = sb; }
    }
}

bool readFile(const std::string& path, std::string& out)
{
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? (size_t)n : 0);
    const size_t got = out.empty() ? 0 : std::fread(&out[0], 1, out.size(), f);
    std::fclose(f);
    return got == out.size();
}

} // namespace

struct okrt_obj { std::vector<LoadedMesh> meshes; };

namespace {

int buildAsset(Parser& ps, float scale, okrt_obj** out)
{
    okrt_obj* asset = new okrt_obj();
    /* ObjFileImporter::createNodes: objects in file order, their meshes in creation order, empty meshes dropped */
    for (const ObjObject& obj : ps.objects)
        for (int mi : obj.meshes) {
            const ObjMesh& m = ps.meshes[mi];
            if (m.faces.empty()) continue;
            /* createVertexArray: one vertex per face corner */
            std::vector<V3f> pos, nor, uv;
            std::vector<std::vector<uint32_t>> faces;
            for (const Face& f : m.faces) {
                std::vector<uint32_t> idx;
                for (const Corner& c : f.corners) {
                    idx.push_back((uint32_t)pos.size());
                    pos.push_back(ps.positions[c.v]);
                    nor.push_back(c.vn >= 0 ? ps.normals[c.vn] : V3f{0, 0, 0});
                    uv.push_back(c.vt >= 0 ? ps.uvs[c.vt] : V3f{0, 0, 0});
                }
                faces.push_back(std::move(idx));
            }
            if (!m.hasNormals) { delete asset; okrt::setError("okrt_obj_load: mesh '" + m.name + "' has no normals (the reference dereferences aiMesh::mNormals unconditionally, Importer.cpp:38)"); return OKRT_ERR_INVALID_ARGUMENT; }
            /* ConvertToLeftHanded = MakeLeftHanded + FlipUVs + FlipWindingOrder */
            for (V3f& p : pos) p.z = -p.z;
            for (V3f& n : nor) n.z = -n.z;
            if (m.hasUvs) for (V3f& t : uv) t.y = 1.0f - t.y;
            for (auto& idx : faces) std::reverse(idx.begin(), idx.end());
            /* Triangulate */
            std::vector<uint32_t> tris;
            for (const auto& idx : faces) triangulate(pos, idx, tris);
            /* CalcTangentSpace needs normals and UV channel 0 */
            std::vector<V3f> tang, bitang;
            const bool hasTangents = m.hasUvs && !tris.empty();
            if (hasTangents) calcTangents(pos, nor, uv, tris, tang, bitang);

            /* Importer.cpp:15-56 */
            LoadedMesh lm;
            lm.name = m.name;
            for (int k = 0; k < 3; k++) { lm.box.min[k] = FLT_MAX; lm.box.max[k] = -FLT_MAX; }
            const size_t nt = tris.size() / 3;
            lm.pos.resize(nt); lm.info.resize(nt);
            for (size_t t = 0; t < nt; t++)
                for (int j = 0; j < 3; j++) {
                    const uint32_t i = tris[t * 3 + j];
                    const float p[3] = {pos[i].x * scale, pos[i].y * scale, pos[i].z * scale};
                    okrt_vertex_info& vi = lm.info[t].vertexInfo[j];
                    for (int k = 0; k < 3; k++) {
                        lm.pos[t].position[j][k] = p[k];
                        lm.box.min[k] = std::min(p[k], lm.box.min[k]);
                        lm.box.max[k] = std::max(p[k], lm.box.max[k]);
                    }
                    vi.normal[0] = nor[i].x; vi.normal[1] = nor[i].y; vi.normal[2] = nor[i].z;
                    vi.uv[0] = m.hasUvs ? uv[i].x : 0.f; vi.uv[1] = m.hasUvs ? uv[i].y : 0.f;
                    const V3f tg = hasTangents ? tang[i] : V3f{0, 0, 0}, bt = hasTangents ? bitang[i] : V3f{0, 0, 0};
                    vi.tangent[0] = tg.x; vi.tangent[1] = tg.y; vi.tangent[2] = tg.z;
                    vi.bitangent[0] = bt.x; vi.bitangent[1] = bt.y; vi.bitangent[2] = bt.z;
                }
            if (m.material >= 0 && m.material < (int)ps.materials.size()) {
                const ObjMaterial& mat = ps.materials[m.material];
                for (int s = 0; s < 5; s++) lm.tex[s] = mat.tex[s];
                if (lm.tex[OKRT_OBJ_TEX_NORMAL].empty()) lm.tex[OKRT_OBJ_TEX_NORMAL] = mat.texDisplacement;     /* Importer.cpp:111-114 */
            }
            asset->meshes.push_back(std::move(lm));
        }
    *out = asset;
    return OKRT_OK;
}

} // namespace

extern "C" {

int okrt_obj_load_memory(const char* objText, size_t objLen, const char* mtlText, size_t mtlLen, float scale, okrt_obj** out)
{
    if (!out || (!objText && objLen)) { okrt::setError("okrt_obj_load_memory: null argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    *out = nullptr;
    Parser ps;
    if (mtlText && mtlLen) ps.parseMtl(mtlText, mtlText + mtlLen);
    ps.parse(objText, objText + objLen);
    return buildAsset(ps, scale, out);
}

int okrt_obj_load(const char* path, float scale, okrt_obj** out)
{
    if (!out || !path) { okrt::setError("okrt_obj_load: null argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    *out = nullptr;
    std::string text;
    if (!readFile(path, text)) { okrt::setError(std::string("okrt_obj_load: cannot read ") + path); return OKRT_ERR_INVALID_ARGUMENT; }
    /* materials first: usemtl looks them up by name.  mtllib lines are collected by a first pass over the text. */
    Parser scan;
    {
        const char* p = text.data(); const char* e = p + text.size();
        while (p < e) {
            const char* le = p; while (le < e && *le != '\n') le++;
            const char* s = skipSpace(p, le); const char* t = tokenEnd(s, le);
            if (std::string(s, t) == "mtllib") scan.mtlLibs.push_back(restOfLine(t, le));
            p = le < e ? le + 1 : e;
        }
    }
    Parser ps;
    std::string dir(path);
    const size_t slash = dir.find_last_of("/\\");
    dir = slash == std::string::npos ? std::string() : dir.substr(0, slash + 1);
    for (const std::string& lib : scan.mtlLibs) {
        std::string mtl;
        if (readFile(dir + lib, mtl)) ps.parseMtl(mtl.data(), mtl.data() + mtl.size());     /* a missing .mtl is not an error for assimp either */
    }
    ps.parse(text.data(), text.data() + text.size());
    return buildAsset(ps, scale, out);
}

void okrt_obj_free(okrt_obj* a) { delete a; }
uint32_t okrt_obj_num_meshes(const okrt_obj* a) { return a ? (uint32_t)a->meshes.size() : 0u; }

int okrt_obj_mesh(const okrt_obj* a, uint32_t i, okrt_mesh_view* view, const char** name)
{
    if (!a || i >= a->meshes.size() || !view) { okrt::setError("okrt_obj_mesh: bad argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    const LoadedMesh& m = a->meshes[i];
    view->trianglesPos = m.pos.data(); view->trianglesInfo = m.info.data();
    view->numTriangles = (uint32_t)m.pos.size(); view->boundingBox = m.box;
    if (name) *name = m.name.c_str();
    return OKRT_OK;
}

const char* okrt_obj_texture_path(const okrt_obj* a, uint32_t mesh, uint32_t slot)
{
    if (!a || mesh >= a->meshes.size() || slot >= 5u) return "";
    return a->meshes[mesh].tex[slot].c_str();
}

} // extern "C"
