/*
 * Stand-in for MVE's <mve/image.h>: the part of mve::FloatImage libs/geom/volume.h and volume_io.h use (dimensions,
 * raw byte access, create).  TEST INFRASTRUCTURE ONLY (oracle/_ref).
 */
#ifndef ORACLE_REF_SHIM_MVE_IMAGE_H
#define ORACLE_REF_SHIM_MVE_IMAGE_H

#include <cstddef>
#include <memory>
#include <vector>

namespace mve {

class FloatImage {
public:
    typedef std::shared_ptr<FloatImage> Ptr;
    typedef std::shared_ptr<FloatImage const> ConstPtr;

    static Ptr create(int width, int height, int channels) { return Ptr(new FloatImage(width, height, channels)); }
    int width(void) const { return w; }
    int height(void) const { return h; }
    int channels(void) const { return c; }
    float & at(int index) { return data[index]; }
    float const & at(int index) const { return data[index]; }
    char * get_byte_pointer(void) { return reinterpret_cast<char *>(data.data()); }
    char const * get_byte_pointer(void) const { return reinterpret_cast<char const *>(data.data()); }
    std::size_t get_byte_size(void) const { return data.size() * sizeof(float); }

private:
    FloatImage(int w, int h, int c) : w(w), h(h), c(c), data((std::size_t)w * h * c, 0.0f) {}
    int w, h, c;
    std::vector<float> data;
};

}  // namespace mve

#endif
