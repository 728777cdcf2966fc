// goxel.cpp — Goxel text model loader (host side of the scene path).
//
// Restates what the reference's Sprite.load does per line (reference data.py:271-289):
//     params = line.strip().split(" ")
//     if params[0].isdigit() and params[1].isdigit() and params[2].isdigit() and params[3] in materials:
//         post = size.x - int(params[0]), int(params[2]), int(params[1]);  voxels[post] = materials[params[3]]
// i.e. fields are separated by exactly one space (an empty field is not a number), header / comment lines fail
// the isdigit test, unknown colour keys are dropped, Y and Z are swapped, X is mirrored around size.x (so file
// column 0 lands on x == size.x, outside the drawn box), later lines override earlier ones.  Voxels are returned in
// file order so that the caller's scatter keeps "last one wins".  `.gz` files are read through zlib like
// gzip.open (data.py:278).  Where the interpreter would raise (a line with fewer than four fields whose first
// fields are numbers: IndexError) this loader reports an error instead of guessing.
#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include "../../include/voxtrace.h"

namespace {

inline bool py_space(unsigned char c) { return c == ' ' || (c >= 9 && c <= 13) || (c >= 0x1c && c <= 0x1f); }

struct Field { const char *p; size_t n; };

// str.isdigit() for ASCII text: non-empty, digits only
inline bool is_digits(const Field &f) {
    if (f.n == 0) return false;
    for (size_t i = 0; i < f.n; i++) if (f.p[i] < '0' || f.p[i] > '9') return false;
    return true;
}

inline bool to_int(const Field &f, int64_t *out) {        // int() of a digit string, bounded to what a grid can hold
    int64_t v = 0;
    for (size_t i = 0; i < f.n; i++) {
        v = v * 10 + (f.p[i] - '0');
        if (v > 0x7fffffff) return false;
    }
    *out = v;
    return true;
}

struct Parser {
    int32_t size_x;
    std::vector<std::string> keys;
    std::vector<int32_t> xyz;
    std::vector<uint8_t> mat;
    int64_t line_no = 0;
    std::string err;

    bool line(const char *b, const char *e) {
        line_no++;
        while (b < e && py_space((unsigned char)*b)) b++;
        while (e > b && py_space((unsigned char)e[-1])) e--;
        Field f[4];
        int nf = 0;
        const char *p = b;
        // str.split(" "): every single space separates; an empty line yields one empty field
        for (;;) {
            const char *q = (const char *)memchr(p, ' ', (size_t)(e - p));
            const char *end = q ? q : e;
            if (nf < 4) { f[nf].p = p; f[nf].n = (size_t)(end - p); }
            nf++;
            if (!q || nf >= 4) break;
            p = q + 1;
        }
        // short-circuit order of the reference's condition; indexing past the list raises in Python
        for (int k = 0; k < 3; k++) {
            if (k >= nf) return fail_short();
            if (!is_digits(f[k])) return true;
        }
        if (nf < 4) return fail_short();
        int key = -1;
        for (size_t k = 0; k < keys.size(); k++)
            if (keys[k].size() == f[3].n && memcmp(keys[k].data(), f[3].p, f[3].n) == 0) { key = (int)k; break; }
        if (key < 0) return true;
        int64_t x, y, z;
        if (!to_int(f[0], &x) || !to_int(f[1], &y) || !to_int(f[2], &z)) {
            err = "line " + std::to_string(line_no) + ": coordinate does not fit 32 bits";
            return false;
        }
        xyz.push_back((int32_t)((int64_t)size_x - x)); xyz.push_back((int32_t)z); xyz.push_back((int32_t)y);
        mat.push_back((uint8_t)(key + 1));
        return true;
    }
    bool fail_short() {
        err = "line " + std::to_string(line_no) + ": fewer than four fields after numeric ones (the reference raises IndexError here)";
        return false;
    }
    // whole text with universal newlines: \n, \r\n and \r all end a line
    bool text(const char *b, const char *e) {
        const char *p = b;
        while (p < e) {
            const char *q = p;
            while (q < e && *q != '\n' && *q != '\r') q++;
            if (!line(p, q)) return false;
            if (q + 1 < e && *q == '\r' && q[1] == '\n') q++;
            p = q < e ? q + 1 : e;
        }
        return true;
    }
};

bool ends_with(const char *s, const char *suf) {
    size_t a = strlen(s), b = strlen(suf);
    return a >= b && strcmp(s + a - b, suf) == 0;
}

void set_err(char *err, int err_len, const std::string &m) {
    if (err && err_len > 0) { strncpy(err, m.c_str(), (size_t)err_len - 1); err[err_len - 1] = 0; }
}

}  // namespace

extern "C" void vt_free(void *p) { free(p); }

extern "C" int vt_goxel_parse(const char *path, int32_t size_x, const char *const *keys, int n_keys, int32_t **out_xyz,
                              uint8_t **out_mat, int64_t *out_n, char *err, int err_len) {
    if (!path || !out_xyz || !out_mat || !out_n || n_keys < 0 || n_keys > VT_MAX_MATERIALS || (n_keys && !keys)) {
        set_err(err, err_len, "vt_goxel_parse: bad arguments");
        return 1;
    }
    *out_xyz = nullptr; *out_mat = nullptr; *out_n = 0;
    Parser ps;
    ps.size_x = size_x;
    for (int k = 0; k < n_keys; k++) ps.keys.emplace_back(keys[k] ? keys[k] : "");
    std::vector<char> buf;
    const char *b = nullptr, *e = nullptr;
    void *map = nullptr; size_t map_len = 0;
    if (ends_with(path, ".gz")) {                                   // data.py:277-278
        gzFile g = gzopen(path, "rb");
        if (!g) { set_err(err, err_len, std::string("cannot open ") + path + ": " + strerror(errno)); return 2; }
        gzbuffer(g, 1 << 20);
        char chunk[1 << 16];
        int n;
        while ((n = gzread(g, chunk, sizeof chunk)) > 0) buf.insert(buf.end(), chunk, chunk + n);
        int zerr = 0;
        const char *zmsg = gzerror(g, &zerr);
        bool bad = n < 0 || (zerr != Z_OK && zerr != Z_STREAM_END);
        std::string m = bad ? std::string("gzip error in ") + path + ": " + (zmsg ? zmsg : "?") : "";
        gzclose(g);
        if (bad) { set_err(err, err_len, m); return 2; }
        b = buf.data(); e = b + buf.size();
    } else if (ends_with(path, ".txt")) {                           // data.py:275-276
        int fd = open(path, O_RDONLY);
        if (fd < 0) { set_err(err, err_len, std::string("cannot open ") + path + ": " + strerror(errno)); return 2; }
        struct stat st;
        if (fstat(fd, &st) != 0) { close(fd); set_err(err, err_len, "fstat failed"); return 2; }
        map_len = (size_t)st.st_size;
        if (map_len) {
            map = mmap(nullptr, map_len, PROT_READ, MAP_PRIVATE, fd, 0);
            if (map == MAP_FAILED) { close(fd); set_err(err, err_len, "mmap failed"); return 2; }
            madvise(map, map_len, MADV_SEQUENTIAL);
        }
        close(fd);
        b = (const char *)map; e = b + map_len;
    } else {
        // the reference prints a warning and returns for any other extension (data.py:279-281)
        set_err(err, err_len, std::string("cannot open sprite ") + path + ": extension must be .txt or .gz");
        return 2;
    }
    ps.xyz.reserve((size_t)(e - b) / 4);
    bool ok = ps.text(b, e);
    if (map) munmap(map, map_len);
    if (!ok) { set_err(err, err_len, std::string(path) + ": " + ps.err); return 3; }
    int64_t n = (int64_t)ps.mat.size();
    if (n) {
        *out_xyz = (int32_t *)malloc((size_t)n * 3 * sizeof(int32_t));
        *out_mat = (uint8_t *)malloc((size_t)n);
        if (!*out_xyz || !*out_mat) { free(*out_xyz); free(*out_mat); *out_xyz = nullptr; *out_mat = nullptr; set_err(err, err_len, "out of memory"); return 4; }
        memcpy(*out_xyz, ps.xyz.data(), (size_t)n * 3 * sizeof(int32_t));
        memcpy(*out_mat, ps.mat.data(), (size_t)n);
    }
    *out_n = n;
    return 0;
}
