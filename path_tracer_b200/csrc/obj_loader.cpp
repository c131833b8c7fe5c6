// obj_loader.cpp — ptb_load_obj / ptb_load_obj_text: the reference's only on-disk input format (load_obj, blas.rs:44-131) as a
// product entry point.  Host code, as upstream: one pass over the file image, no per-line allocations.
//
// Rules kept (each cited where it is applied): `v` / `vn` / `f` lines only; 1-based indices with a dummy element 0; `vn` normalised
// on load; every face token needs three '/'-separated fields; an index that is not an unsigned decimal is read as signed and added
// to the current array length (wrapping, as release-mode Rust); fan triangulation (refs[0], refs[i], refs[i+1]); normal index 0 ->
// the UN-normalised face normal cross(p1 - p0, p2 - p0) of that fan triangle.  Where the reference panics (unwrap on a bad number,
// index out of bounds, an empty line's tokens[0]) this returns PTB_ERR_INVALID with the line number.
#include <charconv>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/ptb.h"
#include "exact_math.h"

namespace {

thread_local std::string g_obj_err;

// Rust's split_whitespace on the ASCII subset of White_Space (the generators and exporters this loader meets write ASCII)
inline bool is_ws(unsigned char c) { return c == ' ' || (c >= 0x09 && c <= 0x0d); }

struct Tok
{
    const char* p;
    size_t n;
};

// str::parse::<f32>(): [+-]? (inf | infinity | nan | digits [. digits] [e [+-] digits]) with at least one digit in the mantissa;
// correctly rounded.  std::from_chars is correctly rounded too but takes no leading '+' and would read a prefix: both handled here.
bool parse_f32(Tok t, float* out)
{
    const char* p = t.p;
    const char* e = t.p + t.n;
    if (p == e) return false;
    bool neg = false;
    if (*p == '+' || *p == '-')
    {
        neg = *p == '-';
        ++p;
    }
    if (p == e) return false;
    auto ieq = [&](const char* w) {
        const size_t n = std::strlen(w);
        if ((size_t)(e - p) != n) return false;
        for (size_t i = 0; i < n; ++i)
            if ((p[i] | 0x20) != w[i]) return false;
        return true;
    };
    if (ieq("inf") || ieq("infinity"))
    {
        *out = neg ? -ex::inf() : ex::inf();
        return true;
    }
    if (ieq("nan"))
    {
        *out = ex::from_bits(0x7fc00000u);
        return true;
    }
    // grammar check (from_chars would accept a numeric prefix of the token and hex is not Rust syntax anyway)
    const char* q = p;
    size_t digits = 0;
    while (q < e && *q >= '0' && *q <= '9') { ++q; ++digits; }
    if (q < e && *q == '.')
    {
        ++q;
        while (q < e && *q >= '0' && *q <= '9') { ++q; ++digits; }
    }
    if (digits == 0) return false;
    if (q < e && (*q == 'e' || *q == 'E'))
    {
        ++q;
        if (q < e && (*q == '+' || *q == '-')) ++q;
        size_t ed = 0;
        while (q < e && *q >= '0' && *q <= '9') { ++q; ++ed; }
        if (ed == 0) return false;
    }
    if (q != e) return false;
    // from_chars wants no '+' after the exponent marker either ("1e+5" is fine for it) and no leading '.'-less quirks: it takes ".5" and "5."
    float v = 0.0f;
    const auto r = std::from_chars(p, e, v, std::chars_format::general);
    if (r.ec == std::errc::result_out_of_range)
    {
        // Rust rounds overflow to +-inf and underflow to +-0; from_chars reports both as out of range: decide by the exponent sign via strtod
        std::string s(p, e);
        const double d = std::strtod(s.c_str(), nullptr);
        v = d > 1.0 ? ex::inf() : 0.0f;
    }
    else if (r.ec != std::errc() || r.ptr != e) return false;
    *out = neg ? -v : v;
    return true;
}

// indices[k].parse::<usize>().unwrap_or_else(|_| len + indices[k].parse::<isize>().unwrap() as usize)   blas.rs:86-92
bool parse_index(Tok t, size_t len, size_t* out)
{
    const char* p = t.p;
    const char* e = t.p + t.n;
    if (p == e) return false;
    bool neg = false, sign = false;
    if (*p == '+' || *p == '-')
    {
        neg = *p == '-';
        sign = true;
        ++p;
    }
    (void)sign;
    if (p == e) return false;
    unsigned long long v = 0;
    for (const char* q = p; q < e; ++q)
    {
        if (*q < '0' || *q > '9') return false;
        if (v > (~0ull - 9) / 10) return false;  // overflow: both parses fail upstream
        v = v * 10 + (unsigned)(*q - '0');
    }
    if (!neg) *out = (size_t)v;                       // usize::from_str takes an optional '+'
    else *out = len + (size_t)(0ull - v);             // isize as usize, wrapping add ("-0" -> len)
    return true;
}

struct Loader
{
    std::vector<float> pos{0.0f, 0.0f, 0.0f}, nrm{0.0f, 0.0f, 0.0f};  // element 0 = Vec3A::ZERO (blas.rs:46-47)
    std::vector<float> out;
    std::vector<Tok> tok;
    std::vector<size_t> rv, rn;

    int fail(size_t line, const std::string& m)
    {
        g_obj_err = "OBJ line " + std::to_string(line) + ": " + m;
        return PTB_ERR_INVALID;
    }

    int run(const char* text, size_t len)
    {
        const char* p = text;
        const char* end = text + len;
        size_t line_no = 0;
        while (p < end)
        {
            const char* nl = (const char*)std::memchr(p, '\n', (size_t)(end - p));
            const char* le = nl ? nl : end;  // BufRead::lines: split at '\n' (a trailing "\r" is whitespace to split_whitespace anyway)
            ++line_no;
            tok.clear();
            for (const char* q = p; q < le;)
            {
                while (q < le && is_ws((unsigned char)*q)) ++q;
                if (q == le) break;
                const char* s = q;
                while (q < le && !is_ws((unsigned char)*q)) ++q;
                tok.push_back({s, (size_t)(q - s)});
            }
            p = nl ? nl + 1 : end;
            if (tok.empty()) return fail(line_no, "empty line (the reference indexes tokens[0], blas.rs:58)");
            const Tok k = tok[0];
            const bool is_v = k.n == 1 && k.p[0] == 'v', is_vn = k.n == 2 && k.p[0] == 'v' && k.p[1] == 'n', is_f = k.n == 1 && k.p[0] == 'f';
            if (is_v || is_vn)
            {
                if (tok.size() < 4) return fail(line_no, "fewer than three coordinates");
                float c[3];
                for (int i = 0; i < 3; ++i)
                    if (!parse_f32(tok[1 + i], &c[i])) return fail(line_no, "not an f32: '" + std::string(tok[1 + i].p, tok[1 + i].n) + "'");
                if (is_v) pos.insert(pos.end(), c, c + 3);
                else
                {
                    const ex::V3 n = ex::normalize(ex::v3(c[0], c[1], c[2]));  // blas.rs:75
                    nrm.push_back(n.x); nrm.push_back(n.y); nrm.push_back(n.z);
                }
            }
            else if (is_f)
            {
                rv.clear();
                rn.clear();
                const size_t np = pos.size() / 3, nn = nrm.size() / 3;
                for (size_t t = 1; t < tok.size(); ++t)
                {
                    // token.split('/'): fields 0 and 2 are used, field 1 (vt) is ignored, a missing field 2 is an index panic (blas.rs:90)
                    const char* s = tok[t].p;
                    const char* e = s + tok[t].n;
                    Tok f[3];
                    int nf = 0;
                    const char* fs = s;
                    for (const char* q = s;; ++q)
                        if (q == e || *q == '/')
                        {
                            if (nf < 3) f[nf] = {fs, (size_t)(q - fs)};
                            ++nf;
                            fs = q + 1;
                            if (q == e) break;
                        }
                    if (nf < 3) return fail(line_no, "face token '" + std::string(s, e) + "' has no normal field (v/vt/vn or v//vn required, blas.rs:90)");
                    size_t v, vn;
                    if (!parse_index(f[0], np, &v) || !parse_index(f[2], nn, &vn)) return fail(line_no, "bad index in face token '" + std::string(s, e) + "'");
                    rv.push_back(v);
                    rn.push_back(vn);
                }
                if (rv.empty()) return fail(line_no, "face without vertices (refs.len() - 1 underflows, blas.rs:97)");
                for (size_t i = 1; i + 1 < rv.size(); ++i)  // 1..(refs.len() - 1): empty for one or two references
                {
                    const size_t iv[3] = {rv[0], rv[i], rv[i + 1]}, in[3] = {rn[0], rn[i], rn[i + 1]};
                    for (int c = 0; c < 3; ++c)
                        if (iv[c] >= np || (in[c] != 0 && in[c] >= nn)) return fail(line_no, "index out of range");
                    const float* P0 = &pos[iv[0] * 3];
                    const float* P1 = &pos[iv[1] * 3];
                    const float* P2 = &pos[iv[2] * 3];
                    bool have_face = false;
                    ex::V3 fn = ex::v3(0, 0, 0);
                    for (int c = 0; c < 3; ++c)
                    {
                        const float* P = &pos[iv[c] * 3];
                        out.push_back(P[0]); out.push_back(P[1]); out.push_back(P[2]);
                        if (in[c] != 0)
                        {
                            const float* N = &nrm[in[c] * 3];
                            out.push_back(N[0]); out.push_back(N[1]); out.push_back(N[2]);
                        }
                        else
                        {
                            if (!have_face)  // blas.rs:113-115, un-normalised
                            {
                                fn = ex::cross(ex::sub(ex::v3(P1[0], P1[1], P1[2]), ex::v3(P0[0], P0[1], P0[2])), ex::sub(ex::v3(P2[0], P2[1], P2[2]), ex::v3(P0[0], P0[1], P0[2])));
                                have_face = true;
                            }
                            out.push_back(fn.x); out.push_back(fn.y); out.push_back(fn.z);
                        }
                    }
                }
            }
            // comments, vt, g, usemtl, unknown keywords: skipped (blas.rs:122-126)
        }
        return PTB_OK;
    }
};

int finish(Loader& L, int rc, float** verts, uint64_t* ntri)
{
    if (rc != PTB_OK) return rc;
    const size_t n = L.out.size();
    float* buf = (float*)std::malloc(std::max<size_t>(n, 1) * sizeof(float));
    if (!buf)
    {
        g_obj_err = "out of memory";
        return PTB_ERR_OOM;
    }
    if (n) std::memcpy(buf, L.out.data(), n * sizeof(float));
    *verts = buf;
    *ntri = n / 18;
    return PTB_OK;
}

}  // namespace

extern "C" {

int ptb_load_obj_text(const char* text, size_t len, float** verts, uint64_t* ntri)
{
    if (!verts || !ntri || (!text && len)) return PTB_ERR_INVALID;
    *verts = nullptr;
    *ntri = 0;
    Loader L;
    return finish(L, L.run(text, len), verts, ntri);
}

int ptb_load_obj(const char* path, float** verts, uint64_t* ntri)
{
    if (!path || !verts || !ntri) return PTB_ERR_INVALID;
    *verts = nullptr;
    *ntri = 0;
    std::FILE* f = std::fopen(path, "rb");
    if (!f)
    {
        g_obj_err = std::string("cannot open '") + path + "' (File::open(path).unwrap(), blas.rs:51)";
        return PTB_ERR_INVALID;
    }
    std::string text;
    char buf[1 << 16];
    size_t n;
    while ((n = std::fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, n);
    std::fclose(f);
    Loader L;
    return finish(L, L.run(text.data(), text.size()), verts, ntri);
}

const char* ptb_load_obj_error(void) { return g_obj_err.c_str(); }

void ptb_free(void* p) { std::free(p); }

}  // extern "C"
