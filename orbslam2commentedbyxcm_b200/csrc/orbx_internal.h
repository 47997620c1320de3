// orbx_internal.h -- internal (non-ABI) bridge between the extractor and matcher translation units of liborbx.so.
#pragma once
#include <cstdint>
#include "../../include/orbx.h"

namespace orbx {

// Device view of one frame's image pyramid inside an extractor handle's arena: interior[l] points at pixel (0,0) of
// level l (the 19-px REFLECT_101 frame lies before / around it), rows are pitch[l] bytes apart.
struct PyramidView {
    int nlevels;
    int device;
    const uint8_t* interior[ORBX_MAX_LEVELS];
    int pitch[ORBX_MAX_LEVELS];
    int w[ORBX_MAX_LEVELS], h[ORBX_MAX_LEVELS];
    float scale[ORBX_MAX_LEVELS], invScale[ORBX_MAX_LEVELS];
};

// Fills `out` for frame `frame` of the last extracted batch; ORBX_ERR_BAD_ARG when nothing was extracted yet.
int pyramid_view(orbx_handle* h, int frame, PyramidView* out);

}  // namespace orbx
