// sib_model.cpp -- model blob parsing and validation.
// The blob replaces model/saved_model.pb (InferenceBolt.java:49): one self-contained little-endian file,
//   "SIBM" | u32 version=1 | u32 in_h,in_w,in_c | u32 n_layers | u32 n_weights | u32 rsvd
//   n_layers x { u32 kind,a,b,c,d,flags,w_off,b_off }   (offsets in floats into the weight section)
//   n_weights x f32    (conv filters HWIO, dense kernels [in,out] -- TensorFlow layouts)
#include <cstring>

#include "sib_internal.hpp"

namespace sib {

static thread_local std::string g_err;
void set_error(const std::string& s) { g_err = s; }
const std::string& get_error() { return g_err; }

int parse_model(const void* blob, size_t n, ModelDesc* m, std::string* err) {
    auto fail = [&](const char* why) {
        if (err) *err = std::string("model blob: ") + why;
        return SIB_EINVAL;
    };
    if (!blob || n < 32) return fail("too short");
    if (std::memcmp(blob, "SIBM", 4) != 0) {
        // not a blob: the reference's own artefact, model/saved_model.pb (InferenceBolt.java:49), is accepted as is
        std::vector<uint8_t> converted;
        std::string ierr;
        if (import_pb(blob, n, &converted, &ierr) != SIB_OK) {
            if (err) *err = "model: bad magic, neither a SIBM blob nor an importable saved_model.pb (" + ierr + ")";
            return SIB_EINVAL;
        }
        return parse_model(converted.data(), converted.size(), m, err);
    }
    const uint32_t* w = static_cast<const uint32_t*>(blob);
    if (w[1] != 1) return fail("unsupported version");
    *m = ModelDesc{};
    m->in_h = w[2]; m->in_w = w[3]; m->in_c = w[4]; m->n_layers = w[5]; m->n_weights = w[6];
    if (m->n_layers == 0 || m->n_layers > (uint32_t)kMaxLayers) return fail("layer count out of range");
    if (m->in_h == 0 || m->in_w == 0 || m->in_c == 0 || m->in_h > 4096 || m->in_w > 4096 || m->in_c > 64)
        return fail("input shape out of range");
    size_t need = 32 + (size_t)m->n_layers * 32 + (size_t)m->n_weights * 4;
    if (n < need) return fail("truncated");
    std::memcpy(m->layer, w + 8, (size_t)m->n_layers * 32);
    const float* wt = reinterpret_cast<const float*>(w + 8 + m->n_layers * 8);
    m->weights.assign(wt, wt + m->n_weights);

    uint32_t h = m->in_h, wd = m->in_w, c = m->in_c, feat = h * wd * c;
    bool flat = false;
    m->img_elems = feat;
    m->max_store_act = feat;
    for (uint32_t i = 0; i < m->n_layers; i++) {
        const LayerDesc& L = m->layer[i];
        if (L.kind == kKindConv) {
            if (flat) return fail("conv after dense");
            if (L.c != c) return fail("conv cin mismatch");
            if (L.a == 0 || L.b == 0 || h < L.a || wd < L.b) return fail("conv kernel larger than map");
            uint32_t oh = h - L.a + 1, ow = wd - L.b + 1;
            if (oh < 2 || ow < 2) return fail("map too small to pool");
            if ((size_t)L.w_off + (size_t)L.a * L.b * L.c * L.d > m->n_weights || (size_t)L.b_off + L.d > m->n_weights)
                return fail("conv weights out of range");
            m->flops_per_image += 2ll * oh * ow * L.d * (L.a * L.b * L.c);
            h = oh / 2; wd = ow / 2; c = L.d; feat = h * wd * c;
        } else if (L.kind == kKindDense) {
            flat = true;
            if (L.a != feat) return fail("dense input size mismatch");
            if ((size_t)L.w_off + (size_t)L.a * L.b > m->n_weights || (size_t)L.b_off + L.b > m->n_weights)
                return fail("dense weights out of range");
            m->flops_per_image += 2ll * L.a * L.b;
            feat = L.b;
        } else {
            return fail("unknown layer kind");
        }
        if (feat > m->max_store_act) m->max_store_act = feat;
    }
    if (!flat) return fail("no dense head");
    if (feat == 0 || feat > (uint32_t)kMaxClasses) return fail("class count out of range");
    m->n_classes = feat;

    // preset recognition (shapes only) -- the tensor-core kernels are specialised on these
    auto is = [&](uint32_t H, uint32_t W, uint32_t C, uint32_t c1, uint32_t c2, std::initializer_list<uint32_t> fc) {
        if (m->in_h != H || m->in_w != W || m->in_c != C) return false;
        if (m->n_layers != 2 + fc.size()) return false;
        const LayerDesc* L = m->layer;
        if (L[0].kind != kKindConv || L[0].a != 5 || L[0].b != 5 || L[0].d != c1) return false;
        if (L[1].kind != kKindConv || L[1].a != 5 || L[1].b != 5 || L[1].d != c2) return false;
        size_t k = 2;
        for (uint32_t f : fc) {
            if (L[k].kind != kKindDense || L[k].b != f) return false;
            bool last = (k + 1 == m->n_layers);
            if (((L[k].flags & 1u) != 0) == last) return false;
            k++;
        }
        return true;
    };
    if (is(28, 28, 1, 6, 16, {120, 84, 10})) m->preset = 0;
    else if (is(32, 32, 3, 6, 16, {120, 84, 10})) m->preset = 1;
    else if (is(28, 28, 1, 16, 32, {64, 10})) m->preset = 2;
    else if (is(32, 32, 3, 16, 32, {64, 10})) m->preset = 3;
    return SIB_OK;
}

}  // namespace sib
