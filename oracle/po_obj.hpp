// po_obj.hpp — CPU restatement of load_obj (src/tlas/tlas_bvh/blas.rs:44-131), statement by statement.
// TEST INFRASTRUCTURE (see po_math.hpp): the checker for ptb_load_obj; never linked into libptb.so.  PARITY UNPINNED: the
// reference ships no OBJ fixtures; Rust's str::parse::<f32> is correctly rounded, as glibc strtof is.
#pragma once
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "po_math.hpp"

namespace po {

struct ObjError
{
    std::string msg;
};

inline std::vector<std::string> split_whitespace(const std::string& line)  // str::split_whitespace, ASCII White_Space
{
    std::vector<std::string> t;
    size_t i = 0;
    auto ws = [](unsigned char c) { return c == ' ' || (c >= 9 && c <= 13); };
    while (i < line.size())
    {
        while (i < line.size() && ws((unsigned char)line[i])) ++i;
        size_t j = i;
        while (j < line.size() && !ws((unsigned char)line[j])) ++j;
        if (j > i) t.push_back(line.substr(i, j - i));
        i = j;
    }
    return t;
}

inline float obj_parse_f32(const std::string& s)  // tokens[k].parse::<f32>().unwrap()   :62-64, :71-73
{
    // Rust's grammar: optional sign, then inf / infinity / nan (any case) or a decimal with optional exponent; no hex, no prefix parse
    size_t i = 0;
    if (i < s.size() && (s[i] == '+' || s[i] == '-')) ++i;
    std::string body = s.substr(i);
    std::string low = body;
    for (auto& c : low) c = (char)std::tolower((unsigned char)c);
    bool ok = low == "inf" || low == "infinity" || low == "nan";
    if (!ok)
    {
        size_t k = 0, digits = 0;
        while (k < body.size() && std::isdigit((unsigned char)body[k])) { ++k; ++digits; }
        if (k < body.size() && body[k] == '.')
        {
            ++k;
            while (k < body.size() && std::isdigit((unsigned char)body[k])) { ++k; ++digits; }
        }
        ok = digits > 0;
        if (ok && k < body.size() && (body[k] == 'e' || body[k] == 'E'))
        {
            ++k;
            if (k < body.size() && (body[k] == '+' || body[k] == '-')) ++k;
            size_t ed = 0;
            while (k < body.size() && std::isdigit((unsigned char)body[k])) { ++k; ++ed; }
            ok = ed > 0;
        }
        ok = ok && k == body.size();
    }
    if (!ok) throw ObjError{"not an f32: '" + s + "'"};
    return std::strtof(s.c_str(), nullptr);
}

inline size_t obj_parse_index(const std::string& s, size_t len)  // :86-92
{
    auto all_digits = [](const std::string& x, size_t from) {
        if (from >= x.size()) return false;
        for (size_t i = from; i < x.size(); ++i)
            if (!std::isdigit((unsigned char)x[i])) return false;
        return true;
    };
    // parse::<usize>(): optional '+', digits
    if (all_digits(s, (!s.empty() && s[0] == '+') ? 1 : 0)) return (size_t)std::strtoull(s.c_str() + ((s[0] == '+') ? 1 : 0), nullptr, 10);
    // parse::<isize>().unwrap() as usize, added to len (wrapping in release builds)
    if (!s.empty() && s[0] == '-' && all_digits(s, 1)) return len + (size_t)(0ull - std::strtoull(s.c_str() + 1, nullptr, 10));
    throw ObjError{"bad index '" + s + "'"};
}

struct ObjVertex
{
    V3 position, normal;
};

inline std::vector<ObjVertex> load_obj_text(const std::string& text)
{
    std::vector<V3> normals{V3{0, 0, 0}};    // :46
    std::vector<V3> positions{V3{0, 0, 0}};  // :47
    std::vector<ObjVertex> vertices;          // :49
    size_t start = 0, line_no = 0;
    while (start < text.size())
    {
        size_t nl = text.find('\n', start);
        std::string line = text.substr(start, nl == std::string::npos ? std::string::npos : nl - start);
        start = nl == std::string::npos ? text.size() : nl + 1;
        if (!line.empty() && line.back() == '\r') line.pop_back();  // BufRead::lines strips "\r\n"
        ++line_no;
        try
        {
            const std::vector<std::string> tokens = split_whitespace(line);  // :56
            if (tokens.empty()) throw ObjError{"empty line: tokens[0] out of bounds"};  // :58
            const std::string& kw = tokens[0];
            if (kw == "v")  // :60-67
            {
                if (tokens.size() < 4) throw ObjError{"fewer than three coordinates"};
                positions.push_back(V3{obj_parse_f32(tokens[1]), obj_parse_f32(tokens[2]), obj_parse_f32(tokens[3])});
            }
            else if (kw == "vn")  // :68-76
            {
                if (tokens.size() < 4) throw ObjError{"fewer than three coordinates"};
                normals.push_back(normalize(V3{obj_parse_f32(tokens[1]), obj_parse_f32(tokens[2]), obj_parse_f32(tokens[3])}));
            }
            else if (kw == "f")  // :77-120
            {
                struct Ref { size_t vertex, normal; };
                std::vector<Ref> refs;
                for (size_t t = 1; t < tokens.size(); ++t)
                {
                    std::vector<std::string> indices;  // token.split('/')
                    size_t a = 0;
                    for (;;)
                    {
                        size_t b = tokens[t].find('/', a);
                        indices.push_back(tokens[t].substr(a, b == std::string::npos ? std::string::npos : b - a));
                        if (b == std::string::npos) break;
                        a = b + 1;
                    }
                    const size_t v = obj_parse_index(indices[0], positions.size());
                    if (indices.size() < 3) throw ObjError{"indices[2] out of bounds"};  // :90
                    const size_t vn = obj_parse_index(indices[2], normals.size());
                    refs.push_back(Ref{v, vn});
                }
                if (refs.empty()) throw ObjError{"refs.len() - 1 underflows"};  // :97 (wraps in release, then refs[1] is out of bounds)
                for (size_t i = 1; i < refs.size() - 1; ++i)  // :97
                {
                    const Ref p[3] = {refs[0], refs[i], refs[i + 1]};
                    for (const Ref& vr : p)
                    {
                        if (vr.vertex >= positions.size()) throw ObjError{"positions index out of bounds"};
                        V3 normal;
                        if (vr.normal != 0)
                        {
                            if (vr.normal >= normals.size()) throw ObjError{"normals index out of bounds"};
                            normal = normals[vr.normal];
                        }
                        else
                        {
                            if (p[1].vertex >= positions.size() || p[2].vertex >= positions.size() || p[0].vertex >= positions.size()) throw ObjError{"positions index out of bounds"};
                            const V3 u = positions[p[1].vertex] - positions[p[0].vertex];
                            const V3 v = positions[p[2].vertex] - positions[p[0].vertex];
                            normal = cross(u, v);  // :113-115
                        }
                        vertices.push_back(ObjVertex{positions[vr.vertex], normal});
                    }
                }
            }
        }
        catch (ObjError& e)
        {
            throw ObjError{"line " + std::to_string(line_no) + ": " + e.msg};
        }
    }
    return vertices;
}

}  // namespace po
