/*
 * nrrd_read.h -- TEST INFRASTRUCTURE (oracle side).  Minimal NRRD0001..5 reader.
 *
 * The reference loads its density volume with teem's nrrdLoad
 * (cuda_codes/trace_rays_through_density_gradients.h:1676-1690) and then reads
 *   axis[i].size, axis[i].spacing, spaceOrigin[i], data      (:1687-1712).
 * teem is a binary-only dependency (libteem 1.11/1.12, not loadable here), so this
 * restates the published NRRD text-header format for exactly those fields:
 *   type (float/double/short/uchar...), dimension 3, sizes, spacings | space directions,
 *   space origin, endian, encoding raw|gzip, optional detached "data file".
 * Pinned by sample-data/bos/sample-density.nrrd (NRRD0005, float, 64^3, raw).
 */
#ifndef ORACLE_NRRD_READ_H_
#define ORACLE_NRRD_READ_H_

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>
#include <unistd.h>
#include <zlib.h>

namespace onrrd {

struct Volume {
    int nx = 0, ny = 0, nz = 0;
    double spacing[3] = {NAN, NAN, NAN};
    double origin[3] = {0, 0, 0};
    std::vector<float> rho;  // x fastest
    std::string error;
};

static inline std::string trim(const std::string &s) {
    size_t a = s.find_first_not_of(" \t\r\n");
    if (a == std::string::npos) return "";
    size_t b = s.find_last_not_of(" \t\r\n");
    return s.substr(a, b - a + 1);
}

static inline bool parse_vec(const std::string &tok, double out[3]) {
    // "(a,b,c)"
    return sscanf(tok.c_str(), " ( %lf , %lf , %lf )", &out[0], &out[1], &out[2]) == 3;
}

static inline bool load(const char *path, Volume &v) {
    FILE *f = fopen(path, "rb");
    if (!f) { v.error = std::string("cannot open ") + path; return false; }
    std::string type = "float", encoding = "raw", endian = "little", datafile;
    int dim = 0;
    bool have_sizes = false;
    long line_skip = 0, byte_skip = 0;
    char buf[4096];
    bool magic = false;
    while (fgets(buf, sizeof buf, f)) {
        std::string line = trim(buf);
        if (!magic) {
            if (line.compare(0, 4, "NRRD") != 0) { v.error = "not a NRRD file"; fclose(f); return false; }
            magic = true;
            continue;
        }
        if (line.empty()) break;  // blank line ends the header
        if (line[0] == '#') continue;
        size_t c = line.find(':');
        if (c == std::string::npos) continue;
        if (c + 1 < line.size() && line[c + 1] == '=') continue;  // key/value pair
        std::string key = trim(line.substr(0, c)), val = trim(line.substr(c + 1));
        if (key == "type") type = val;
        else if (key == "dimension") dim = atoi(val.c_str());
        else if (key == "sizes") {
            have_sizes = sscanf(val.c_str(), "%d %d %d", &v.nx, &v.ny, &v.nz) == 3;
        } else if (key == "spacings") {
            sscanf(val.c_str(), "%lf %lf %lf", &v.spacing[0], &v.spacing[1], &v.spacing[2]);
        } else if (key == "space directions") {
            // teem leaves axis.spacing NaN when only space directions are given; we
            // take the vector lengths so such files are still usable.
            const char *p = val.c_str();
            for (int a = 0; a < 3; a++) {
                const char *o = strchr(p, '(');
                if (!o) break;
                double d[3];
                if (parse_vec(o, d) && std::isnan(v.spacing[a]))
                    v.spacing[a] = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
                p = strchr(o, ')');
                if (!p) break;
                p++;
            }
        } else if (key == "space origin") parse_vec(val, v.origin);
        else if (key == "endian") endian = val;
        else if (key == "encoding") encoding = val;
        else if (key == "data file" || key == "datafile") datafile = val;
        else if (key == "line skip" || key == "lineskip") line_skip = atol(val.c_str());
        else if (key == "byte skip" || key == "byteskip") byte_skip = atol(val.c_str());
    }
    if (dim != 3 || !have_sizes || v.nx <= 0 || v.ny <= 0 || v.nz <= 0) {
        v.error = "need dimension 3 with sizes"; fclose(f); return false;
    }
    if (endian != "little") { v.error = "only little-endian supported"; fclose(f); return false; }
    size_t esz = 0;
    enum { F32, F64, I16, U16, U8, I8, I32, U32 } kind = F32;
    if (type == "float") { kind = F32; esz = 4; }
    else if (type == "double") { kind = F64; esz = 8; }
    else if (type == "short" || type == "int16" || type == "signed short") { kind = I16; esz = 2; }
    else if (type == "ushort" || type == "uint16" || type == "unsigned short") { kind = U16; esz = 2; }
    else if (type == "uchar" || type == "uint8" || type == "unsigned char") { kind = U8; esz = 1; }
    else if (type == "signed char" || type == "int8") { kind = I8; esz = 1; }
    else if (type == "int" || type == "int32" || type == "signed int") { kind = I32; esz = 4; }
    else if (type == "uint" || type == "uint32" || type == "unsigned int") { kind = U32; esz = 4; }
    else { v.error = "unsupported type " + type; fclose(f); return false; }

    size_t n = (size_t)v.nx * v.ny * v.nz;
    std::vector<unsigned char> raw(n * esz);
    FILE *df = f;
    if (!datafile.empty()) {
        std::string p = datafile;
        if (p[0] != '/') {
            std::string base(path);
            size_t s = base.find_last_of('/');
            p = (s == std::string::npos ? std::string("") : base.substr(0, s + 1)) + datafile;
        }
        df = fopen(p.c_str(), "rb");
        if (!df) { v.error = "cannot open data file " + p; fclose(f); return false; }
    }
    for (long i = 0; i < line_skip; i++) { if (!fgets(buf, sizeof buf, df)) break; }
    bool ok = true;
    if (encoding == "raw") {
        if (byte_skip > 0) fseek(df, byte_skip, SEEK_CUR);
        else if (byte_skip == -1) fseek(df, -(long)(n * esz), SEEK_END);
        ok = fread(raw.data(), 1, n * esz, df) == n * esz;
    } else if (encoding == "gzip" || encoding == "gz") {
        long pos = ftell(df);
        int fd = dup(fileno(df));
        lseek(fd, pos, SEEK_SET);
        gzFile g = gzdopen(fd, "rb");
        if (!g) ok = false;
        else {
            std::vector<unsigned char> skip((size_t)(byte_skip > 0 ? byte_skip : 0));
            if (!skip.empty()) gzread(g, skip.data(), (unsigned)skip.size());
            size_t got = 0;
            while (got < n * esz) {
                unsigned want = (unsigned)std::min<size_t>(n * esz - got, 1u << 30);
                int r = gzread(g, raw.data() + got, want);
                if (r <= 0) break;
                got += (size_t)r;
            }
            ok = got == n * esz;
            gzclose(g);
        }
    } else { v.error = "unsupported encoding " + encoding; ok = false; }
    if (df != f) fclose(df);
    fclose(f);
    if (!ok) { if (v.error.empty()) v.error = "short read of voxel data"; return false; }

    v.rho.resize(n);
    for (size_t i = 0; i < n; i++) {
        const unsigned char *p = raw.data() + i * esz;
        float val;
        switch (kind) {
            case F32: memcpy(&val, p, 4); break;
            case F64: { double d; memcpy(&d, p, 8); val = (float)d; } break;
            case I16: { int16_t d; memcpy(&d, p, 2); val = (float)d; } break;
            case U16: { uint16_t d; memcpy(&d, p, 2); val = (float)d; } break;
            case U8: val = (float)*p; break;
            case I8: val = (float)*(const signed char *)p; break;
            case I32: { int32_t d; memcpy(&d, p, 4); val = (float)d; } break;
            default: { uint32_t d; memcpy(&d, p, 4); val = (float)d; } break;
        }
        v.rho[i] = val;
    }
    return true;
}

}  // namespace onrrd
#endif
