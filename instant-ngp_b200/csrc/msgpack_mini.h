// MessagePack encoder / decoder over ngp::Json, as much of the format as nlohmann::json::to_msgpack / from_msgpack emit and
// accept for instant-ngp snapshots (reference: src/testbed.cu:4873-4936 save_snapshot, :246-273 load_network_config;
// spec: github.com/msgpack/msgpack/blob/master/spec.md).  Integers take the smallest encoding, floats are written as float32
// when that is exact and float64 otherwise (nlohmann's rule), binary blobs as bin 8/16/32, maps with string keys.
#pragma once
#include "json_mini.h"
#include <cstring>

namespace ngp {

namespace msgpack_detail {
inline void put_be(std::string& o, uint64_t v, int bytes) { for (int k = bytes - 1; k >= 0; --k) o.push_back((char)((v >> (8 * k)) & 0xFF)); }
inline void put_len(std::string& o, size_t n, uint8_t fix_base, size_t fix_max, uint8_t t8, uint8_t t16, uint8_t t32) {
    if (fix_max && n <= fix_max) { o.push_back((char)(fix_base | n)); return; }
    if (t8 && n <= 0xFF) { o.push_back((char)t8); put_be(o, n, 1); return; }
    if (n <= 0xFFFF) { o.push_back((char)t16); put_be(o, n, 2); return; }
    if (n > 0xFFFFFFFFull) throw std::runtime_error("msgpack: container larger than 4 GiB");
    o.push_back((char)t32); put_be(o, n, 4);
}
}  // namespace msgpack_detail

inline void msgpack_encode(const Json& j, std::string& o) {
    using namespace msgpack_detail;
    switch (j.type) {
        case Json::Null: o.push_back((char)0xC0); break;
        case Json::Bool: o.push_back((char)(j.b ? 0xC3 : 0xC2)); break;
        case Json::Number: {
            const double v = j.num;
            if (!j.is_float && v == (double)(int64_t)v && v >= -9.2e18 && v <= 1.8e19) {
                if (v >= 0) {
                    const uint64_t u = (uint64_t)v;
                    if (u < 128) o.push_back((char)u);
                    else if (u <= 0xFF) { o.push_back((char)0xCC); put_be(o, u, 1); }
                    else if (u <= 0xFFFF) { o.push_back((char)0xCD); put_be(o, u, 2); }
                    else if (u <= 0xFFFFFFFFull) { o.push_back((char)0xCE); put_be(o, u, 4); }
                    else { o.push_back((char)0xCF); put_be(o, u, 8); }
                } else {
                    const int64_t i = (int64_t)v;
                    if (i >= -32) o.push_back((char)(int8_t)i);
                    else if (i >= -128) { o.push_back((char)0xD0); put_be(o, (uint64_t)i, 1); }
                    else if (i >= -32768) { o.push_back((char)0xD1); put_be(o, (uint64_t)i, 2); }
                    else if (i >= -2147483648ll) { o.push_back((char)0xD2); put_be(o, (uint64_t)i, 4); }
                    else { o.push_back((char)0xD3); put_be(o, (uint64_t)i, 8); }
                }
            } else if ((double)(float)v == v) {
                const float f = (float)v; uint32_t u; memcpy(&u, &f, 4); o.push_back((char)0xCA); put_be(o, u, 4);
            } else { uint64_t u; memcpy(&u, &v, 8); o.push_back((char)0xCB); put_be(o, u, 8); }
            break;
        }
        case Json::String: put_len(o, j.str.size(), 0xA0, 31, 0xD9, 0xDA, 0xDB); o += j.str; break;
        case Json::Binary: put_len(o, j.str.size(), 0, 0, 0xC4, 0xC5, 0xC6); o += j.str; break;
        case Json::Array: put_len(o, j.arr.size(), 0x90, 15, 0, 0xDC, 0xDD); for (const Json& e : j.arr) msgpack_encode(e, o); break;
        case Json::Object:
            put_len(o, j.obj.size(), 0x80, 15, 0, 0xDE, 0xDF);
            for (const auto& kv : j.obj) { put_len(o, kv.first.size(), 0xA0, 31, 0xD9, 0xDA, 0xDB); o += kv.first; msgpack_encode(kv.second, o); }
            break;
    }
}

namespace msgpack_detail {
struct Reader {
    const uint8_t* p; size_t n, i = 0;
    [[noreturn]] void fail(const char* what) const { throw std::runtime_error(std::string("msgpack parse error at offset ") + std::to_string(i) + ": " + what); }
    void need(size_t k) const { if (k > n - i) fail("unexpected end of input"); }
    uint64_t be(int bytes) { need(bytes); uint64_t v = 0; for (int k = 0; k < bytes; ++k) v = (v << 8) | p[i++]; return v; }
    std::string bytes(size_t k) { need(k); std::string s((const char*)p + i, k); i += k; return s; }
    Json value(int depth) {
        if (depth > 64) fail("nesting too deep");
        need(1);
        const uint8_t t = p[i++];
        Json j;
        if (t < 0x80) return Json::integer(t);
        if (t >= 0xE0) return Json::integer((int8_t)t);
        if (t >= 0xA0 && t <= 0xBF) return Json::string(bytes(t & 31));
        if (t >= 0x90 && t <= 0x9F) return array(t & 15, depth);
        if (t >= 0x80 && t <= 0x8F) return map(t & 15, depth);
        switch (t) {
            case 0xC0: return j;
            case 0xC2: return Json::boolean(false);
            case 0xC3: return Json::boolean(true);
            case 0xC4: case 0xC5: case 0xC6: { const size_t k = be(1 << (t - 0xC4)); j.type = Json::Binary; j.str = bytes(k); return j; }
            case 0xCA: { const uint32_t u = (uint32_t)be(4); float f; memcpy(&f, &u, 4); return Json::number(f, true); }
            case 0xCB: { const uint64_t u = be(8); double d; memcpy(&d, &u, 8); return Json::number(d, true); }
            case 0xCC: return Json::integer((int64_t)be(1));
            case 0xCD: return Json::integer((int64_t)be(2));
            case 0xCE: return Json::integer((int64_t)be(4));
            case 0xCF: return Json::number((double)be(8), false);
            case 0xD0: return Json::integer((int8_t)be(1));
            case 0xD1: return Json::integer((int16_t)be(2));
            case 0xD2: return Json::integer((int32_t)be(4));
            case 0xD3: return Json::integer((int64_t)be(8));
            case 0xD9: case 0xDA: case 0xDB: return Json::string(bytes(be(1 << (t - 0xD9))));
            case 0xDC: return array(be(2), depth);
            case 0xDD: return array(be(4), depth);
            case 0xDE: return map(be(2), depth);
            case 0xDF: return map(be(4), depth);
            default: fail("unsupported type byte (ext types do not occur in snapshots)");
        }
    }
    Json array(size_t k, int depth) { Json j; j.type = Json::Array; if (k > n - i) fail("array longer than input"); j.arr.reserve(k); for (size_t e = 0; e < k; ++e) j.arr.push_back(value(depth + 1)); return j; }
    Json map(size_t k, int depth) {
        Json j; j.type = Json::Object;
        for (size_t e = 0; e < k; ++e) {
            Json key = value(depth + 1);
            if (key.type != Json::String) fail("map key is not a string");
            j.obj[key.str] = value(depth + 1);
        }
        return j;
    }
};
}  // namespace msgpack_detail

inline Json msgpack_decode(const void* data, size_t size) {
    msgpack_detail::Reader r{(const uint8_t*)data, size};
    Json j = r.value(0);
    if (r.i != size) r.fail("trailing bytes");
    return j;
}

}  // namespace ngp
