// TEST INFRASTRUCTURE ONLY.  C entry points around the reference's own metadata.bin writer and reader
// (/root/reference/src/metadata.cpp, compiled unchanged by oracle/build_ref.sh into oracle/_ref/libref_metadata.so),
// used to pin the sidecar format: tests/golden/make_metadata_golden.py writes the committed fixture through them.
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include "metadata.h"

extern "C" {

// names: n zero-terminated strings back to back
int refmeta_write(const char* path, int n, const int32_t* label, const int32_t* x, const int32_t* y, const int32_t* w,
                  const int32_t* h, const char* names, uint32_t image_w, uint32_t image_h) {
    try {
        std::vector<SliceMetadata> m(n);
        const char* p = names;
        for (int i = 0; i < n; ++i) {
            m[i].label = label[i]; m[i].x = x[i]; m[i].y = y[i]; m[i].width = w[i]; m[i].height = h[i];
            m[i].filename = p;
            p += m[i].filename.size() + 1;
        }
        write_metadata_binary(m, path, image_w, image_h);
        return 0;
    } catch (...) { return 1; }
}

// returns the record count (or -1); fills up to cap records, names back to back into names_out (cap_names bytes)
int refmeta_read(const char* path, int cap, int32_t* label, int32_t* x, int32_t* y, int32_t* w, int32_t* h, char* names_out,
                 int cap_names, uint32_t* image_w, uint32_t* image_h) {
    try {
        std::vector<SliceMetadata> m = read_metadata_binary(path, *image_w, *image_h);
        int used = 0;
        for (int i = 0; i < (int)m.size() && i < cap; ++i) {
            label[i] = m[i].label; x[i] = m[i].x; y[i] = m[i].y; w[i] = m[i].width; h[i] = m[i].height;
            const int len = (int)m[i].filename.size() + 1;
            if (used + len > cap_names) return -1;
            std::memcpy(names_out + used, m[i].filename.c_str(), len);
            used += len;
        }
        return (int)m.size();
    } catch (...) { return -1; }
}

}
