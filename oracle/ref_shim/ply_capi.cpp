// C entry points over the reference's own PLYLoader (src/ply_loader.{h,cpp}, compiled
// unmodified where it lies under /root/reference). TEST INFRASTRUCTURE ONLY: pins the
// host-side loader restatement and the synthetic PLY writer.
#include "ply_loader.h"
#include <cstring>

extern "C" {

// Returns a handle (nullptr on failure); *n = number of PointData records (62 floats each).
void* refply_load(const char* path, unsigned* n) {
    auto* pts = new std::vector<PointData>();
    if (!PLYLoader::load(path, *pts)) { delete pts; if (n) *n = 0; return nullptr; }
    if (n) *n = (unsigned)pts->size();
    return pts;
}
const float* refply_data(void* h) { return reinterpret_cast<const float*>(static_cast<std::vector<PointData>*>(h)->data()); }
unsigned refply_record_bytes() { return (unsigned)sizeof(PointData); }
void refply_free(void* h) { delete static_cast<std::vector<PointData>*>(h); }

}  // extern "C"
