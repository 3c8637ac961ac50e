// flatten.h — host-side glue between the reference's scene/camera types and the POD
// arrays of the C ABI (include/prt.h). Header-only; compiled against the UNMODIFIED
// reference headers (-I<reference>/src -I<reference>/external/glm ...).
//
// What it walks (paths relative to /root/reference):
//   Scene::GetPrimitives()            src/core/scene.h:29-32
//   Primitive{Shape, Material, Transform}  src/core/primitive.h:7-12
//   ShapeHandle / MaterialHandle tags  src/core/shape.h:10-15, src/core/material_handle.h:13-19
//   Circle::getRadius, Quad::GetWidth/GetHeight  src/core/shape.h:25,39-40
//   material getters                   src/core/material.h:38,64-65,102,129
//   Camera getters + basis             src/core/camera.h:10-16,96-99
// The reference uploads 160-byte Primitives holding host pointers and two 4x4 matrices
// (cuda_wavefront/soa.cpp:37-114); here every transform is baked into world-space records
// and no host pointer ever reaches the device.
#pragma once
#include <core/scene.h>
#include <core/camera.h>
#include <core/mesh.h>
#include <cmath>
#include <cstring>
#include <map>
#include <vector>
#include "prt.h"

namespace prt_host {

struct FlatScene {
    std::vector<prt_sphere> spheres;
    std::vector<prt_quad> quads;
    std::vector<float> tri_verts, tri_normals;   // 9 floats per triangle
    std::vector<int32_t> tri_material;
    std::vector<prt_material> materials;
    int tri_prim_base = 0;                       // triangles must be contiguous in the list
    bool ok = true;
    std::string error;
};

// layout-compatible view of the reference's Triangle (members are private, shape.h:78-81)
struct TrianglePod { glm::vec3 v[3]; glm::vec3 n[3]; glm::vec2 uv[3]; };
static_assert(sizeof(TrianglePod) == sizeof(Triangle), "Triangle layout changed");

inline int FlattenMaterial(const MaterialHandle& h, std::map<const void*, int>& seen, std::vector<prt_material>& out) {
    if (!h.IsValid()) return -1;
    auto it = seen.find(h.Ptr);
    if (it != seen.end()) return it->second;
    prt_material m{};
    m.type = (int32_t)h.Tag();
    if (auto* l = h.As<MatType::LAMBERTIAN>()) { auto a = l->GetAlbedo(); m.r = a.x; m.g = a.y; m.b = a.z; }
    else if (auto* me = h.As<MatType::METAL>()) { auto a = me->GetAlbedo(); m.r = a.x; m.g = a.y; m.b = a.z; m.param = me->GetRoughness(); }
    else if (auto* d = h.As<MatType::DIELECTRIC>()) { m.r = m.g = m.b = 1.0f; m.param = d->GetRefractionIndex(); }
    else if (auto* e = h.As<MatType::EMISSIVE>()) { auto a = e->GetEmission(); m.r = a.x; m.g = a.y; m.b = a.z; }
    int idx = (int)out.size();
    out.push_back(m);
    seen[h.Ptr] = idx;
    return idx;
}

// Mesh -> scene glue: the reference's Scene has no triangle pool (shape.h:107-126, scene.h:41) although its
// Mesh loader exists (mesh.h:8-21). Appends every face of `mesh`, translated by `offset`, as world-space triangles with
// material `mat` AFTER the primitives already flattened (so prim_id order = scene primitives, then meshes in AddMesh order).
// Vertex normals as the Mesh holds them (mesh.h:13); a mesh without normals gets flat face normals.
inline void AppendMesh(FlatScene& fs, const Mesh& mesh, const prt_material& mat, const glm::vec3& offset) {
    const auto& V = mesh.GetVertices(); const auto& N = mesh.GetNormals(); const auto& I = mesh.GetIndices();
    const int mi = (int)fs.materials.size();
    fs.materials.push_back(mat);
    const bool has_n = N.size() == V.size() && !N.empty();
    for (size_t f = 0; f + 2 < I.size(); f += 3) {
        glm::vec3 p[3], n[3];
        for (int k = 0; k < 3; ++k) p[k] = V[I[f + k]] + offset;
        if (has_n) for (int k = 0; k < 3; ++k) n[k] = N[I[f + k]];
        else { const glm::vec3 fn = glm::normalize(glm::cross(p[1] - p[0], p[2] - p[0])); n[0] = n[1] = n[2] = fn; }
        for (int k = 0; k < 3; ++k) {
            fs.tri_verts.push_back(p[k].x); fs.tri_verts.push_back(p[k].y); fs.tri_verts.push_back(p[k].z);
            fs.tri_normals.push_back(n[k].x); fs.tri_normals.push_back(n[k].y); fs.tri_normals.push_back(n[k].z);
        }
        fs.tri_material.push_back(mi);
    }
}

inline FlatScene FlattenPrimitives(const std::vector<Primitive>& prims) {
    FlatScene fs;
    std::map<const void*, int> seen;
    int first_tri = -1, last_tri = -1;
    for (int i = 0; i < (int)prims.size(); ++i) {
        const Primitive& p = prims[i];
        const glm::mat4& M = p.Transform.GetMat();
        const glm::mat4& Mi = p.Transform.GetInvMat();
        glm::vec3 c0(M[0]), c1(M[1]), c2(M[2]), tr(M[3]);
        float s0 = glm::length(c0), s2 = glm::length(c2);
        int mat = FlattenMaterial(p.Material, seen, fs.materials);
        if (auto* ci = p.Shape.As<ShapeType::CIRCLE>()) {
            prt_sphere s{};
            s.cx = tr.x; s.cy = tr.y; s.cz = tr.z;
            s.radius = ci->getRadius() * s0;
            s.tmin = kShapeRayTMin * s0;
            s.material = mat; s.prim_id = i;
            fs.spheres.push_back(s);
        } else if (auto* q = p.Shape.As<ShapeType::QUAD>()) {
            prt_quad r{};
            glm::vec3 u = c0 / s0, v = c2 / s2;
            glm::vec3 n = TransformNormal(Mi, glm::vec3(0.0f, 1.0f, 0.0f));   // what primitive.cpp:39 yields for a quad
            r.cx = tr.x; r.cy = tr.y; r.cz = tr.z; r.half_w = (q->GetWidth() / 2.0f) * s0;
            r.ux = u.x; r.uy = u.y; r.uz = u.z; r.half_h = (q->GetHeight() / 2.0f) * s2;
            r.vx = v.x; r.vy = v.y; r.vz = v.z; r.tmin = kShapeRayTMin * glm::length(c1);
            r.nx = n.x; r.ny = n.y; r.nz = n.z;
            r.material = mat; r.prim_id = i;
            fs.quads.push_back(r);
        } else if (auto* t = p.Shape.As<ShapeType::TRIANGLE>()) {
            TrianglePod pod;
            std::memcpy(&pod, t, sizeof pod);
            if (first_tri < 0) first_tri = i;
            if (last_tri >= 0 && i != last_tri + 1) { fs.ok = false; fs.error = "triangles must be contiguous in the primitive list"; }
            last_tri = i;
            for (int k = 0; k < 3; ++k) {
                glm::vec3 w = TransformPoint(M, pod.v[k]);
                glm::vec3 nn = glm::mat3(glm::transpose(Mi)) * pod.n[k];
                fs.tri_verts.push_back(w.x); fs.tri_verts.push_back(w.y); fs.tri_verts.push_back(w.z);
                fs.tri_normals.push_back(nn.x); fs.tri_normals.push_back(nn.y); fs.tri_normals.push_back(nn.z);
            }
            fs.tri_material.push_back(mat);
        } else {
            fs.ok = false; fs.error = "unknown shape tag";
        }
    }
    fs.tri_prim_base = first_tri < 0 ? (int)prims.size() : first_tri;
    return fs;
}

inline FlatScene FlattenScene(const Scene& scene) { return FlattenPrimitives(scene.GetPrimitives()); }

inline prt_camera FlattenCamera(const Camera& cam) {
    prt_camera k{};
    glm::vec3 pos = cam.GetPosition(), front = cam.GetViewDir();
    // m_Right / m_Up are private: recompute exactly as the constructor does (camera.h:13-15)
    glm::vec3 right = glm::normalize(glm::cross(front, glm::vec3(0.0f, 1.0f, 0.0f)));
    glm::vec3 up = glm::normalize(glm::cross(right, front));
    for (int i = 0; i < 3; ++i) { k.pos[i] = pos[i]; k.front[i] = front[i]; k.right[i] = right[i]; k.up[i] = up[i]; }
    float tanFovY = tan(0.5f);                          // camera.h:111
    k.tan_half_fov = tanFovY;
    k.width = (int32_t)cam.GetWidth(); k.height = (int32_t)cam.GetHeight();
    return k;
}

}  // namespace prt_host
