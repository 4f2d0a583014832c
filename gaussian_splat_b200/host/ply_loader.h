// Host-side mirror of the reference loader interface (src/ply_loader.h:7-33): same struct, same
// static entry point, same results on the files the reference reads (binary 3DGS PLYs).
#pragma once
#include <string>
#include <vector>

// 62 floats = 248 bytes, field order as src/ply_loader.h:7-17.
struct PointData {
    float x = 0, y = 0, z = 0;
    float nx = 0, ny = 0, nz = 0;
    float r = 0, g = 0, b = 0;                       // clamped 0.5 + C0 * f_dc
    float opacity = 1.0f;                            // sigmoid(stored logit)
    float scale_x = 0.01f, scale_y = 0.01f, scale_z = 0.01f;   // exp(stored log)
    float rot_0 = 1, rot_1 = 0, rot_2 = 0, rot_3 = 0;          // w x y z, not normalised
    float sh_rest[45] = {};
};
static_assert(sizeof(PointData) == 248, "PointData must stay 62 packed floats");

class PLYLoader {
public:
    // Returns false (and prints to stderr) when the file cannot be opened, the header is not a PLY
    // vertex table, or no point was read. Behaviour notes (kept on purpose):
    //  * every property is read as a 4-byte little-endian float, whatever its declared type;
    //  * an all-zero f_dc triple is left as colour (0,0,0), not converted to 0.5 grey;
    //  * ASCII files yield exactly `element vertex` points (the reference yields 2x: a resize
    //    followed by push_back, src/ply_loader.cpp:52,199 — deliberately not reproduced).
    static bool load(const std::string& filepath, std::vector<PointData>& points);

    // --- addition (no counterpart in the reference): the file's vertex table left UNDECODED, for the GPU
    // decode entry splat_upload_ply_rows (include/splat_cuda.h). rows = n * stride floats exactly as load()
    // would feed them to its per-row decode (binary body, or the parsed numbers of an ASCII file);
    // slot_dst[j] = PointData float index of the j-th property, -1 = ignored.
    struct Rows {
        std::vector<float> rows;
        size_t n = 0;
        unsigned stride = 0;
        std::vector<int> slot_dst;
    };
    static bool loadRows(const std::string& filepath, Rows& out);
};
