// sib_pbimport.cpp -- TensorFlow SavedModel / frozen GraphDef -> model blob.
//
// The reference ships its network as model/saved_model.pb, copies ONLY that file out of the jar and loads it with
// SavedModelBundle.load("./", "serve") (InferenceBolt.java:49-58), so every weight must be a Const inside the graph
// (no variables/ directory travels). It feeds "input:0" and fetches "output/Softmax:0" (InferenceBolt.java:82-83).
// This file reads that artefact without TensorFlow or libprotobuf: a ~100-line protobuf wire reader, the handful of
// message fields that matter (field numbers from tensorflow/core/{protobuf/saved_model,protobuf/meta_graph,
// framework/graph,framework/node_def,framework/attr_value,framework/tensor,framework/tensor_shape}.proto), and a
// backwards walk from the Softmax node that accepts exactly the model family of SURVEY.md D1:
//   Placeholder -> [Conv2D(VALID, stride 1, NHWC) -> BiasAdd -> Relu -> MaxPool(2x2 / 2)] x k -> Reshape
//               -> [MatMul -> BiasAdd/Add -> Relu] x m -> MatMul -> BiasAdd/Add -> Softmax
// (Identity nodes anywhere, Relu and MaxPool in either order). Anything else is refused with a message naming the
// node, never approximated. TF layouts are the blob's layouts (HWIO filters, [in,out] kernels), so weights are
// copied, not permuted. PARITY UNPINNED: no TensorFlow exists on this box and the reference ships no model
// (SURVEY F3); tests/test_pbimport.py builds the protobuf bytes with an independent Python encoder.
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "sib_internal.hpp"

namespace sib {
namespace {

struct View {
    const uint8_t* p = nullptr;
    size_t n = 0;
    std::string str() const { return std::string(reinterpret_cast<const char*>(p), n); }
};

// protobuf wire format: (field << 3 | type) varint, then varint (0) | 8 bytes (1) | length + bytes (2) | 4 bytes (5)
struct PbReader {
    const uint8_t* p;
    const uint8_t* end;
    bool bad = false;
    explicit PbReader(View v) : p(v.p), end(v.p + v.n) {}
    bool varint(uint64_t* out) {
        uint64_t v = 0;
        for (int shift = 0; shift < 64 && p < end; shift += 7) {
            const uint8_t b = *p++;
            v |= (uint64_t)(b & 0x7f) << shift;
            if (!(b & 0x80)) { *out = v; return true; }
        }
        bad = true;
        return false;
    }
    // next field; scalar types land in `num`, length-delimited in `bytes`
    bool next(uint32_t* field, uint32_t* type, uint64_t* num, View* bytes) {
        if (bad || p >= end) return false;
        uint64_t key;
        if (!varint(&key)) return false;
        *field = (uint32_t)(key >> 3);
        *type = (uint32_t)(key & 7);
        *num = 0;
        *bytes = View{};
        switch (*type) {
            case 0: return varint(num);
            case 1: if (end - p < 8) break; std::memcpy(num, p, 8); p += 8; return true;
            case 5: if (end - p < 4) break; std::memcpy(num, p, 4); p += 4; return true;
            case 2: {
                uint64_t len;
                if (!varint(&len) || len > (uint64_t)(end - p)) break;
                bytes->p = p; bytes->n = (size_t)len; p += len;
                return true;
            }
            default: break;
        }
        bad = true;
        return false;
    }
};

struct Node {
    std::string name, op;
    std::vector<std::string> inputs;
    std::map<std::string, View> attr;  // key -> AttrValue message bytes
};

struct Tensor {
    std::vector<int64_t> dims;
    std::vector<float> data;
};

View sub(View msg, uint32_t want_field, int index = 0) {  // index-th length-delimited field `want_field` of `msg`
    PbReader r(msg);
    uint32_t f, t; uint64_t v; View b;
    while (r.next(&f, &t, &v, &b))
        if (f == want_field && t == 2 && index-- == 0) return b;
    return View{};
}

bool parse_node(View msg, Node* nd) {
    PbReader r(msg);
    uint32_t f, t; uint64_t v; View b;
    while (r.next(&f, &t, &v, &b)) {
        if (t != 2) continue;
        if (f == 1) nd->name = b.str();
        else if (f == 2) nd->op = b.str();
        else if (f == 3) nd->inputs.push_back(b.str());
        else if (f == 5) {  // map<string, AttrValue> entry: 1 = key, 2 = value
            const View k = sub(b, 1), val = sub(b, 2);
            if (k.p) nd->attr[k.str()] = val;
        }
    }
    return !r.bad;
}

// GraphDef nodes, wherever the caller's bytes start: SavedModel.meta_graphs(2)[0].graph_def(2), or a bare GraphDef
bool collect_nodes(View file, std::vector<Node>* nodes) {
    auto graph_nodes = [&](View g) {
        PbReader r(g);
        uint32_t f, t; uint64_t v; View b;
        std::vector<Node> out;
        while (r.next(&f, &t, &v, &b))
            if (f == 1 && t == 2) {
                Node nd;
                if (!parse_node(b, &nd) || nd.name.empty() || nd.op.empty()) return std::vector<Node>{};
                out.push_back(std::move(nd));
            }
        return r.bad ? std::vector<Node>{} : out;
    };
    const View meta = sub(file, 2);  // SavedModel.meta_graphs
    if (meta.p) {
        const View g = sub(meta, 2);  // MetaGraphDef.graph_def
        if (g.p) *nodes = graph_nodes(g);
    }
    if (nodes->empty()) *nodes = graph_nodes(file);
    return !nodes->empty();
}

bool attr_string(const Node& nd, const char* key, std::string* out) {
    auto it = nd.attr.find(key);
    if (it == nd.attr.end()) return false;
    const View s = sub(it->second, 2);  // AttrValue.s
    if (!s.p) return false;
    *out = s.str();
    return true;
}

bool attr_bool(const Node& nd, const char* key, bool dflt) {
    auto it = nd.attr.find(key);
    if (it == nd.attr.end()) return dflt;
    PbReader r(it->second);
    uint32_t f, t; uint64_t v; View b;
    while (r.next(&f, &t, &v, &b))
        if (f == 5 && t == 0) return v != 0;  // AttrValue.b
    return dflt;
}

bool attr_int_list(const Node& nd, const char* key, std::vector<int64_t>* out) {  // AttrValue.list(1).i(3), packed or not
    auto it = nd.attr.find(key);
    if (it == nd.attr.end()) return false;
    const View list = sub(it->second, 1);
    PbReader r(list);
    uint32_t f, t; uint64_t v; View b;
    while (r.next(&f, &t, &v, &b)) {
        if (f != 3) continue;
        if (t == 0) out->push_back((int64_t)v);
        else if (t == 2) {
            PbReader pr(b);
            uint64_t x;
            while (pr.p < pr.end && pr.varint(&x)) out->push_back((int64_t)x);
        }
    }
    return true;
}

bool shape_dims(View shape, std::vector<int64_t>* dims) {  // TensorShapeProto.dim(2).size(1)
    PbReader r(shape);
    uint32_t f, t; uint64_t v; View b;
    while (r.next(&f, &t, &v, &b)) {
        if (f == 3 && t == 0 && v) return false;  // unknown_rank
        if (f != 2 || t != 2) continue;
        int64_t size = 0;
        PbReader d(b);
        uint32_t f2, t2; uint64_t v2; View b2;
        while (d.next(&f2, &t2, &v2, &b2))
            if (f2 == 1 && t2 == 0) size = (int64_t)v2;
        dims->push_back(size);
    }
    return !r.bad;
}

// Const node -> fp32 tensor. TensorProto: dtype(1) = DT_FLOAT(1), tensor_shape(2), tensor_content(4) | float_val(5)
bool const_tensor(const Node& nd, Tensor* out, std::string* why) {
    auto it = nd.attr.find("value");
    if (it == nd.attr.end()) { *why = "Const '" + nd.name + "' has no value"; return false; }
    const View tp = sub(it->second, 8);  // AttrValue.tensor
    PbReader r(tp);
    uint32_t f, t; uint64_t v; View b;
    uint64_t dtype = 0;
    View content;
    std::vector<float> vals;
    while (r.next(&f, &t, &v, &b)) {
        if (f == 1 && t == 0) dtype = v;
        else if (f == 2 && t == 2) { if (!shape_dims(b, &out->dims)) { *why = "'" + nd.name + "': unknown rank"; return false; } }
        else if (f == 4 && t == 2) content = b;
        else if (f == 5 && t == 5) { float x; uint32_t u = (uint32_t)v; std::memcpy(&x, &u, 4); vals.push_back(x); }
        else if (f == 5 && t == 2) { vals.resize(b.n / 4); if (b.n >= 4) std::memcpy(vals.data(), b.p, b.n / 4 * 4); }  // (1..3 stray bytes: no element, no copy from a null vector)
    }
    if (r.bad || dtype != 1) { *why = "Const '" + nd.name + "' is not a DT_FLOAT tensor"; return false; }
    size_t numel = 1;
    for (int64_t d : out->dims) {
        if (d <= 0 || d > (1 << 24)) { *why = "'" + nd.name + "': bad dimension"; return false; }
        numel *= (size_t)d;
        if (numel > (size_t)1 << 24) { *why = "'" + nd.name + "': tensor too large"; return false; }
    }
    out->data.resize(numel);
    if (content.n == numel * 4) std::memcpy(out->data.data(), content.p, content.n);  // little-endian, as this host
    else if (vals.size() == numel) out->data = vals;
    else if (vals.size() == 1) std::fill(out->data.begin(), out->data.end(), vals[0]);  // TF's splat encoding
    else { *why = "Const '" + nd.name + "': value size does not match its shape"; return false; }
    return true;
}

struct Graph {
    std::vector<Node> nodes;
    std::map<std::string, const Node*> by_name;
    // "name", "name:0", "^name" -> node, looking through Identity
    const Node* producer(const std::string& ref) const {
        std::string nm = ref;
        if (!nm.empty() && nm[0] == '^') return nullptr;  // control dependency
        const size_t colon = nm.rfind(':');
        if (colon != std::string::npos) nm.resize(colon);
        for (int hop = 0; hop < 64; hop++) {
            auto it = by_name.find(nm);
            if (it == by_name.end()) return nullptr;
            const Node* nd = it->second;
            if (nd->op != "Identity" || nd->inputs.empty()) return nd;
            nm = nd->inputs[0];
            const size_t c2 = nm.rfind(':');
            if (c2 != std::string::npos) nm.resize(c2);
        }
        return nullptr;
    }
    const Node* in(const Node* nd, size_t i) const { return i < nd->inputs.size() ? producer(nd->inputs[i]) : nullptr; }
};

struct ConvL { Tensor w, b; };
struct DenseL { Tensor w, b; bool relu; };

bool is_add(const Node* n) { return n && (n->op == "BiasAdd" || n->op == "Add" || n->op == "AddV2"); }

// weights of `nd` (an input of a compute node): must be a frozen constant
bool weights_of(const Graph& g, const Node* user, size_t i, Tensor* t, std::string* why) {
    const Node* nd = g.in(user, i);
    if (!nd) { *why = "input " + std::to_string(i) + " of '" + user->name + "' not found"; return false; }
    if (nd->op != "Const") {
        *why = "'" + nd->name + "' (" + nd->op + ") feeding '" + user->name + "' is not a Const: the graph must be frozen, "
               "the reference copies saved_model.pb without its variables/ (InferenceBolt.java:49-51)";
        return false;
    }
    return const_tensor(*nd, t, why);
}

bool all_equal(const std::vector<int64_t>& v, std::initializer_list<int64_t> want) {
    return v.size() == want.size() && std::equal(v.begin(), v.end(), want.begin());
}

}  // namespace

// Returns SIB_OK and the blob in *blob, or SIB_EINVAL and a message.
int import_pb(const void* bytes, size_t n, std::vector<uint8_t>* blob, std::string* err) {
    auto fail = [&](const std::string& why) {
        if (err) *err = "saved_model.pb import: " + why;
        return SIB_EINVAL;
    };
    if (!bytes || n < 8) return fail("too short");
    Graph g;
    if (!collect_nodes(View{static_cast<const uint8_t*>(bytes), n}, &g.nodes)) return fail("not a SavedModel or GraphDef protobuf");
    for (const Node& nd : g.nodes) g.by_name[nd.name] = &nd;

    // the fetch: "output/Softmax" (InferenceBolt.java:83), else the graph's only Softmax
    const Node* sm = nullptr;
    auto it = g.by_name.find("output/Softmax");
    if (it != g.by_name.end() && it->second->op == "Softmax") sm = it->second;
    if (!sm) {
        int count = 0;
        for (const Node& nd : g.nodes)
            if (nd.op == "Softmax") { sm = &nd; count++; }
        if (count != 1) return fail(count ? "several Softmax nodes and none is named output/Softmax" : "no Softmax node");
    }
    std::vector<DenseL> dense;  // collected output -> input
    std::vector<ConvL> convs;
    std::string why;
    const Node* cur = g.in(sm, 0);
    bool relu_seen = false;
    // A GraphDef is not guaranteed acyclic (a hostile or corrupt file can feed a node its own output): every step of
    // the walk below is counted, and no chain is longer than the node list.
    size_t steps = 0;
    const size_t max_steps = g.nodes.size() + 8;
    // ---- dense head
    for (;;) {
        if (!cur) return fail("dangling input while walking back from '" + sm->name + "'");
        if (++steps > max_steps || dense.size() > (size_t)kMaxLayers) return fail("cycle in the graph at '" + cur->name + "'");
        if (cur->op == "Relu") { if (relu_seen) return fail("two Relu in a row at '" + cur->name + "'"); relu_seen = true; cur = g.in(cur, 0); continue; }
        if (cur->op == "Reshape" || cur->op == "Squeeze" || cur->op == "Flatten") {
            if (relu_seen) return fail("Relu directly after the flatten at '" + cur->name + "'");
            cur = g.in(cur, 0);
            break;
        }
        if (!is_add(cur)) return fail("expected BiasAdd/Add, Relu or Reshape, found '" + cur->name + "' (" + cur->op + ")");
        const Node* mm = g.in(cur, 0);
        size_t bias_at = 1;
        if (!mm || mm->op != "MatMul") { mm = g.in(cur, 1); bias_at = 0; }
        if (!mm || mm->op != "MatMul") return fail("'" + cur->name + "' does not add a bias to a MatMul");
        if (attr_bool(*mm, "transpose_a", false)) return fail("MatMul '" + mm->name + "' has transpose_a");
        DenseL d;
        d.relu = relu_seen;
        relu_seen = false;
        if (!weights_of(g, cur, bias_at, &d.b, &why) || !weights_of(g, mm, 1, &d.w, &why)) return fail(why);
        if (d.w.dims.size() != 2) return fail("MatMul '" + mm->name + "' kernel is not rank 2");
        if (attr_bool(*mm, "transpose_b", false)) {  // stored [out,in]: bring to TF's usual [in,out]
            const int64_t o = d.w.dims[0], i = d.w.dims[1];
            std::vector<float> tr((size_t)o * i);
            for (int64_t a = 0; a < o; a++)
                for (int64_t b = 0; b < i; b++) tr[(size_t)b * o + a] = d.w.data[(size_t)a * i + b];
            d.w.data.swap(tr);
            d.w.dims = {i, o};
        }
        if (d.b.dims.size() != 1 || d.b.dims[0] != d.w.dims[1]) return fail("bias of '" + cur->name + "' does not match its kernel");
        dense.push_back(std::move(d));
        cur = g.in(mm, 0);
    }
    if (dense.empty()) return fail("no dense layer before the Softmax");
    if (dense.front().relu) return fail("Relu between the last dense layer and the Softmax");
    // ---- conv blocks, back to the placeholder
    const Node* ph = nullptr;
    for (;;) {
        if (!cur) return fail("dangling input below the flatten");
        if (++steps > max_steps || convs.size() > (size_t)kMaxLayers) return fail("cycle in the graph at '" + cur->name + "'");
        if (cur->op == "Placeholder") { ph = cur; break; }
        if (cur->op == "Reshape") { cur = g.in(cur, 0); continue; }
        bool relu = false, pool = false;
        while (cur && (cur->op == "Relu" || cur->op == "MaxPool") && ++steps <= max_steps) {
            if (cur->op == "Relu") { if (relu) return fail("two Relu in one block at '" + cur->name + "'"); relu = true; }
            else {
                if (pool) return fail("two MaxPool in one block at '" + cur->name + "'");
                pool = true;
                std::vector<int64_t> ks, st;
                std::string pad, fmt;
                attr_int_list(*cur, "ksize", &ks);
                attr_int_list(*cur, "strides", &st);
                attr_string(*cur, "padding", &pad);
                if (attr_string(*cur, "data_format", &fmt) && fmt != "NHWC") return fail("MaxPool '" + cur->name + "' is not NHWC");
                if (!all_equal(ks, {1, 2, 2, 1}) || !all_equal(st, {1, 2, 2, 1}) || pad != "VALID")
                    return fail("MaxPool '" + cur->name + "' is not 2x2 / stride 2 / VALID");
            }
            cur = g.in(cur, 0);
        }
        if (!relu || !pool) return fail("conv block without Relu and MaxPool" + (cur ? " above '" + cur->name + "'" : std::string()));
        if (!is_add(cur)) return fail("expected BiasAdd under the Relu/MaxPool, found " + (cur ? "'" + cur->name + "' (" + cur->op + ")" : std::string("nothing")));
        const Node* cv = g.in(cur, 0);
        size_t bias_at = 1;
        if (!cv || cv->op != "Conv2D") { cv = g.in(cur, 1); bias_at = 0; }
        if (!cv || cv->op != "Conv2D") return fail("'" + cur->name + "' does not add a bias to a Conv2D");
        std::vector<int64_t> st, dil;
        std::string pad, fmt;
        attr_int_list(*cv, "strides", &st);
        attr_string(*cv, "padding", &pad);
        if (attr_string(*cv, "data_format", &fmt) && fmt != "NHWC") return fail("Conv2D '" + cv->name + "' is not NHWC");
        if (attr_int_list(*cv, "dilations", &dil) && !dil.empty() && !all_equal(dil, {1, 1, 1, 1})) return fail("Conv2D '" + cv->name + "' is dilated");
        if (!all_equal(st, {1, 1, 1, 1}) || pad != "VALID") return fail("Conv2D '" + cv->name + "' is not stride 1 / VALID");
        ConvL c;
        if (!weights_of(g, cur, bias_at, &c.b, &why) || !weights_of(g, cv, 1, &c.w, &why)) return fail(why);
        if (c.w.dims.size() != 4 || c.b.dims.size() != 1 || c.b.dims[0] != c.w.dims[3]) return fail("Conv2D '" + cv->name + "': filter/bias shapes");
        convs.push_back(std::move(c));
        cur = g.in(cv, 0);
    }
    if (convs.empty()) return fail("no Conv2D block");
    if (convs.size() + dense.size() > (size_t)kMaxLayers) return fail("too many layers");
    // ---- input geometry: the placeholder's static shape [-1,H,W,C]; else a square map that fits the first dense layer
    int64_t H = 0, W = 0;
    const int64_t C = convs.back().w.dims[2];
    {
        std::vector<int64_t> dims;
        auto sh = ph->attr.find("shape");
        if (sh != ph->attr.end()) shape_dims(sub(sh->second, 7), &dims);  // AttrValue.shape
        if (dims.size() == 4 && dims[1] > 0 && dims[2] > 0) {
            if (dims[3] > 0 && dims[3] != C) return fail("placeholder channels do not match the first filter");
            H = dims[1]; W = dims[2];
        } else {
            const int64_t feat = dense.back().w.dims[0];
            for (int64_t s = 4; s <= 1024 && !H; s++) {
                int64_t h = s;
                bool ok = true;
                for (size_t k = convs.size(); k-- > 0 && ok;) {
                    h = h - convs[k].w.dims[0] + 1;
                    ok = h >= 2;
                    h /= 2;
                }
                if (ok && h * h * convs.front().w.dims[3] == feat) H = W = s;
            }
            if (!H) return fail("placeholder '" + ph->name + "' has no static shape and no square input fits the dense layer");
        }
    }
    // ---- emit the blob (layout: sib_model.cpp)
    const uint32_t n_layers = (uint32_t)(convs.size() + dense.size());
    std::vector<uint32_t> table;
    std::vector<float> weights;
    auto put = [&](uint32_t kind, uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t flags, const Tensor& w, const Tensor& bias) {
        const uint32_t w_off = (uint32_t)weights.size();
        weights.insert(weights.end(), w.data.begin(), w.data.end());
        const uint32_t b_off = (uint32_t)weights.size();
        weights.insert(weights.end(), bias.data.begin(), bias.data.end());
        for (uint32_t v : {kind, a, b, c, d, flags, w_off, b_off}) table.push_back(v);
    };
    for (size_t k = convs.size(); k-- > 0;) {
        const ConvL& c = convs[k];
        put(kKindConv, (uint32_t)c.w.dims[0], (uint32_t)c.w.dims[1], (uint32_t)c.w.dims[2], (uint32_t)c.w.dims[3], 1u, c.w, c.b);
    }
    for (size_t k = dense.size(); k-- > 0;) {
        const DenseL& d = dense[k];
        put(kKindDense, (uint32_t)d.w.dims[0], (uint32_t)d.w.dims[1], 0u, 0u, d.relu ? 1u : 0u, d.w, d.b);
    }
    const uint32_t head[8] = {0, 1u, (uint32_t)H, (uint32_t)W, (uint32_t)C, n_layers, (uint32_t)weights.size(), 0u};
    blob->resize(32 + table.size() * 4 + weights.size() * 4);
    std::memcpy(blob->data(), head, 32);
    std::memcpy(blob->data(), "SIBM", 4);
    std::memcpy(blob->data() + 32, table.data(), table.size() * 4);
    std::memcpy(blob->data() + 32 + table.size() * 4, weights.data(), weights.size() * 4);
    ModelDesc check;  // shapes must chain (conv cin, dense input size, class count): parse_model is the arbiter
    std::string perr;
    if (parse_model(blob->data(), blob->size(), &check, &perr) != SIB_OK) return fail("graph does not chain: " + perr);
    return SIB_OK;
}

}  // namespace sib
