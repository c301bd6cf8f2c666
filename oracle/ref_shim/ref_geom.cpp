/*
 * TEST INFRASTRUCTURE (oracle/_ref/libref_geom.so): the reference's own libs/geom/sphere.h, libs/util/numpy_io.h and
 * libs/geom/volume_io.h and libs/opti/nelder_mead.h, included from where they lie under $(REFERENCE)/libs and compiled unmodified against the MVE
 * stand-ins of this directory.  Pins the 1272-vertex direction set of the spherical histogram and the bytes of the
 * .npy / VOL writers (uavmvs_b200/formats.py, uavmvs_sphere_vertices).  Nothing in the product loads this.
 */
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <tuple>
#include <vector>

#include "math/vector.h"
#include "geom/sphere.h"
#include "util/numpy_io.h"
#include "geom/volume_io.h"
#include "opti/nelder_mead.h"

extern "C" {

/* generate_sphere_mesh keeps its working vectors in function-local statics and swaps them out (sphere.h:21, 36, 109-110):
 * only the FIRST call of a process is meaningful, exactly as in the apps.  Returns the vertex count, 0 on a later call. */
uint32_t ref_sphere_vertices(float radius, uint32_t subdivisions, float * out, uint32_t capacity)
{
    static bool used = false;
    if (used) return 0;
    used = true;
    mve::TriangleMesh::Ptr mesh = generate_sphere_mesh(radius, subdivisions);
    std::vector<math::Vec3f> const & v = mesh->get_vertices();
    if (out && capacity >= v.size())
        for (std::size_t i = 0; i < v.size(); ++i) for (int k = 0; k < 3; ++k) out[3 * i + k] = v[i][k];
    return (uint32_t)v.size();
}

/* save_numpy_file(v0[, v1[, v2]], filename), libs/util/numpy_io.h:59-102 (evaluate_heuristic.cu:293 passes three) */
int ref_save_numpy_f32(const char * filename, const float * data, uint32_t rows, uint32_t cols)
{
    try {
        std::vector<std::vector<float> > v(rows);
        for (uint32_t r = 0; r < rows; ++r) v[r].assign(data + (std::size_t)r * cols, data + (std::size_t)(r + 1) * cols);
        std::string f(filename);
        if (rows == 1) save_numpy_file(v[0], f);
        else if (rows == 2) save_numpy_file(v[0], v[1], f);
        else if (rows == 3) save_numpy_file(v[0], v[1], v[2], f);
        else return 2;
        return 0;
    } catch (...) { return 1; }
}

int ref_save_numpy_i32(const char * filename, const int32_t * data, uint32_t rows, uint32_t cols)
{
    try {
        std::vector<std::vector<int32_t> > v(rows);
        for (uint32_t r = 0; r < rows; ++r) v[r].assign(data + (std::size_t)r * cols, data + (std::size_t)(r + 1) * cols);
        std::string f(filename);
        if (rows == 1) save_numpy_file(v[0], f);
        else if (rows == 2) save_numpy_file(v[0], v[1], f);
        else return 2;
        return 0;
    } catch (...) { return 1; }
}

/* save_volume<uint>, libs/geom/volume_io.h:22-47: n images of iw x ih x ic floats at the linear indices idx[] */
int ref_save_volume(const char * filename, uint32_t w, uint32_t h, uint32_t d, const float * mn, const float * mx,
    uint32_t n, const uint32_t * idx, int iw, int ih, int ic, const float * data)
{
    try {
        Volume<uint32_t>::Ptr vol = Volume<uint32_t>::create(w, h, d, math::Vec3f(mn[0], mn[1], mn[2]), math::Vec3f(mx[0], mx[1], mx[2]));
        const std::size_t per = (std::size_t)iw * ih * ic;
        for (uint32_t i = 0; i < n; ++i) {
            mve::FloatImage::Ptr img = mve::FloatImage::create(iw, ih, ic);
            std::memcpy(img->get_byte_pointer(), data + per * i, per * sizeof(float));
            vol->at(idx[i]) = img;
        }
        save_volume<uint32_t>(vol, std::string(filename));
        return 0;
    } catch (...) { return 1; }
}

/* load_volume<uint> round trip: returns the number of images and the header; images are copied out in index order */
int ref_load_volume(const char * filename, uint32_t dims[3], float mn[3], float mx[3], uint32_t * n, uint32_t * idx, uint32_t idx_cap,
    float * data, uint64_t data_cap)
{
    try {
        Volume<uint32_t>::Ptr vol = load_volume<uint32_t>(std::string(filename));
        dims[0] = vol->width(); dims[1] = vol->height(); dims[2] = vol->depth();
        for (int k = 0; k < 3; ++k) { mn[k] = vol->minimum()[k]; mx[k] = vol->maximum()[k]; }
        uint32_t count = 0; uint64_t off = 0;
        for (uint32_t i = 0; i < vol->num_positions(); ++i) {
            mve::FloatImage::Ptr img = vol->at(i);
            if (!img) continue;
            const uint64_t per = img->get_byte_size() / sizeof(float);
            if (count < idx_cap && off + per <= data_cap) { idx[count] = i; std::memcpy(data + off, img->get_byte_pointer(), per * sizeof(float)); }
            off += per; ++count;
        }
        *n = count;
        return 0;
    } catch (...) { return 1; }
}

/* One step of the reference's simplex-downhill (libs/opti/nelder_mead.h:29-121) on a Simplex<3> given as 12 floats;
 * the objective is a C callback (the tests evaluate it with the oracle's K5 / K6). */
typedef float (*ref_objective_t)(const float * pos, void * user);
int ref_nelder_mead_step(float * simplex12, ref_objective_t f, void * user, uint32_t * vid, float * value)
{
    Simplex<3> sx;
    for (int i = 0; i < 4; ++i) sx.verts[i] = math::Vec3f(simplex12[3 * i], simplex12[3 * i + 1], simplex12[3 * i + 2]);
    std::function<float(math::Vector<float, 3>)> func = [&](math::Vector<float, 3> p) -> float { return f(*p, user); };
    std::pair<std::size_t, float> r = nelder_mead<3>(&sx, func);
    for (int i = 0; i < 4; ++i) for (int k = 0; k < 3; ++k) simplex12[3 * i + k] = sx.verts[i][k];
    *vid = (uint32_t)r.first; *value = r.second;
    return 0;
}

}  // extern "C"
