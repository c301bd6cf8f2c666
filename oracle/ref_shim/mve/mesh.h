/*
 * Stand-in for MVE's <mve/mesh.h> (nmoehrle/mve, un-vendored dependency of the reference).  TEST INFRASTRUCTURE ONLY:
 * lets the reference's own libs/geom/sphere.h compile unmodified for oracle/_ref.  Only what generate_sphere_mesh touches:
 * the vertex / face / vertex-normal containers, create(), and recalc_normals() (its result is not used by the pin: the
 * histogram kernels consume the vertices only).
 */
#ifndef ORACLE_REF_SHIM_MVE_MESH_H
#define ORACLE_REF_SHIM_MVE_MESH_H

#include <memory>
#include <vector>

#include "math/vector.h"

typedef unsigned int uint;

namespace mve {

class TriangleMesh {
public:
    typedef std::shared_ptr<TriangleMesh> Ptr;
    typedef std::shared_ptr<TriangleMesh const> ConstPtr;
    typedef std::vector<math::Vec3f> VertexList;
    typedef std::vector<unsigned int> FaceList;

    static Ptr create(void) { return Ptr(new TriangleMesh()); }
    VertexList & get_vertices(void) { return vertices; }
    VertexList const & get_vertices(void) const { return vertices; }
    VertexList & get_vertex_normals(void) { return normals; }
    FaceList & get_faces(void) { return faces; }
    FaceList const & get_faces(void) const { return faces; }
    void recalc_normals(bool, bool) {}
    /* touched by parameterize_spherical (sphere.h:117-193), which no app on the path calls */
    std::vector<math::Vec2f> & get_vertex_texcoords(void) { return texcoords; }
    std::vector<float> & get_vertex_confidences(void) { return confidences; }
    std::vector<float> & get_vertex_values(void) { return values; }
    std::vector<math::Vec4f> & get_vertex_colors(void) { return colors; }
    bool has_vertex_confidences(void) const { return !confidences.empty(); }
    bool has_vertex_values(void) const { return !values.empty(); }
    bool has_vertex_colors(void) const { return !colors.empty(); }

private:
    VertexList vertices, normals;
    FaceList faces;
    std::vector<math::Vec2f> texcoords;
    std::vector<float> confidences, values;
    std::vector<math::Vec4f> colors;
};

}  // namespace mve

#endif
