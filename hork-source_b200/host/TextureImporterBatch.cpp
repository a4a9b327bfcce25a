// TextureImporterBatch.cpp -- hk_texture_import: the reference's Tools/TextureImporter (Main.cpp:58-253) as a batch driver
// over the C ABI (include/hkbcn.h), SURVEY.md 8f row f2.
//
// Same flags and meaning as the reference tool for the part of it that is on this path:
//     -s <file> -o <file> -format <name> [-no_alpha] [-alpha_premult] [-mip_edge_mode <m>] [-mip_filter <f>]
// and, new, `-batch <manifest>`: a text file with one such flag set per line (`#` comments), imported as ONE sharded batch
// (hkbcn_import_batch: every texture whole on one GPU, no collective) instead of one process per texture.
// Giving either -mip_* flag turns mip generation on, as in the reference (Main.cpp:124-143); -format takes the
// TexFormat[] names of Hork/Image/Image.cpp:105-118.
//
// -resample <w> <h> [-resample_edge_mode h v] [-resample_filter h v] scale the source first (Main.cpp:145-192 -> ResampleRawImage)
// through hkbcn_resample_raw_image, on the device.
// Out of scope (stays on the CPU in the reference): file decoding through stb_image / tinyexr / webp.
// Sources are raw containers that need no decoder: binary PGM / PPM (P5 / P6, 8 bit), PFM
// (Pf / PF), or `gen:<kind>:<width>:<height>:<seed>` procedural images (kind: albedo, mask, rough, normal, sky).
// Channel mapping follows CreateImage (Image.cpp:1700-1716): BC1/2/3/7 take RGBA8 (alpha 255 when the file has none), BC4
// the first channel, BC5 the first two, BC6H RGBA32F.
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/hkbcn.h"

namespace {

struct Args {
    std::string src, out, format;
    bool no_alpha = false, alpha_premult = false, mips = false;
    int edge = 0, filter = 0;
    // -resample <w> <h> [-resample_edge_mode h v] [-resample_filter h v] (Tools/TextureImporter/Main.cpp:145-192)
    bool resample = false;
    uint32_t rs_w = 0, rs_h = 0;
    int rs_edge_h = 0, rs_edge_v = 0, rs_filter_h = 0, rs_filter_v = 0;
};

struct Pixels {
    uint32_t w = 0, h = 0;
    int ch = 0;      // channels in `u8` / `f32`
    bool is_float = false;
    std::vector<uint8_t> u8;
    std::vector<float> f32;
};

const struct { const char* name; int fmt; } kFormats[] = {
    {"BC1_UNORM", HKBCN_BC1_UNORM}, {"BC1_UNORM_SRGB", HKBCN_BC1_UNORM_SRGB}, {"BC2_UNORM", HKBCN_BC2_UNORM},
    {"BC2_UNORM_SRGB", HKBCN_BC2_UNORM_SRGB}, {"BC3_UNORM", HKBCN_BC3_UNORM}, {"BC3_UNORM_SRGB", HKBCN_BC3_UNORM_SRGB},
    {"BC4_UNORM", HKBCN_BC4_UNORM}, {"BC5_UNORM", HKBCN_BC5_UNORM}, {"BC6H_UFLOAT", HKBCN_BC6H_UFLOAT},
    {"BC6H_SFLOAT", HKBCN_BC6H_SFLOAT}, {"BC7_UNORM", HKBCN_BC7_UNORM}, {"BC7_UNORM_SRGB", HKBCN_BC7_UNORM_SRGB}};

int find_format(const std::string& n)
{
    for (auto& f : kFormats)
        if (n == f.name) return f.fmt;
    return -1;
}
// GetResampleEdgeMode / GetResampleFilter of the reference tool: names -> IMAGE_RESAMPLE_EDGE_MODE / _FILTER (Image.h:330-347)
int edge_mode(const std::string& s) { return s == "clamp" ? 1 : s == "reflect" ? 2 : s == "wrap" ? 3 : s == "zero" ? 4 : 0; }
int filter_of(const std::string& s) { return s == "box" ? 1 : s == "triangle" ? 2 : s == "cubicspline" ? 3 : s == "catmullrom" ? 4 : s == "mitchell" ? 5 : 0; }

bool parse_flags(const std::vector<std::string>& t, Args& a, std::string& err)
{
    for (size_t i = 0; i < t.size(); i++) {
        auto need = [&](int n) { return i + (size_t)n < t.size(); };
        if (t[i] == "-s" && need(1)) a.src = t[++i];
        else if (t[i] == "-o" && need(1)) a.out = t[++i];
        else if (t[i] == "-format" && need(1)) a.format = t[++i];
        else if (t[i] == "-no_alpha") a.no_alpha = true;
        else if (t[i] == "-alpha_premult") a.alpha_premult = true;
        else if (t[i] == "-mip_edge_mode" && need(1)) { a.edge = edge_mode(t[++i]); a.mips = true; }
        else if (t[i] == "-mip_filter" && need(1)) { a.filter = filter_of(t[++i]); a.mips = true; }
        else if (t[i] == "-resample" && need(2)) {
            a.resample = true;
            a.rs_w = (uint32_t)strtoul(t[i + 1].c_str(), nullptr, 10);
            a.rs_h = (uint32_t)strtoul(t[i + 2].c_str(), nullptr, 10);
            i += 2;
            if (a.rs_w == 0 || a.rs_h == 0) { err = "Invalid resample size " + std::to_string(a.rs_w) + " x " + std::to_string(a.rs_h); return false; }
            if (a.rs_w > 4096 || a.rs_h > 4096) { err = "Resulting texture size is too large " + std::to_string(a.rs_w) + " x " + std::to_string(a.rs_h); return false; }
        } else if (t[i] == "-resample_edge_mode" && need(2)) { a.rs_edge_h = edge_mode(t[i + 1]); a.rs_edge_v = edge_mode(t[i + 2]); i += 2; }
        else if (t[i] == "-resample_filter" && need(2)) { a.rs_filter_h = filter_of(t[i + 1]); a.rs_filter_v = filter_of(t[i + 2]); i += 2; }
        else { err = "unknown or incomplete flag '" + t[i] + "'"; return false; }
    }
    if (a.out.empty()) { err = "Output file is not specified. Use -o <filename>"; return false; }
    if (a.src.empty()) { err = "Source file is not specified. Use -s <filename>"; return false; }
    if (find_format(a.format) < 0) { err = "unknown -format '" + a.format + "'"; return false; }
    return true;
}

// ---- sources ----------------------------------------------------------------------------------------------------------
uint32_t lcg(uint32_t& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

bool generate(const std::string& spec, Pixels& p, std::string& err)
{
    // gen:<kind>:<w>:<h>:<seed> -- smooth gradients + value noise, enough to exercise every encoder mode
    std::vector<std::string> f;
    std::stringstream ss(spec);
    for (std::string x; std::getline(ss, x, ':');) f.push_back(x);
    if (f.size() != 5) { err = "expected gen:<kind>:<w>:<h>:<seed>"; return false; }
    p.w = (uint32_t)atoi(f[2].c_str());
    p.h = (uint32_t)atoi(f[3].c_str());
    uint32_t seed = (uint32_t)strtoul(f[4].c_str(), nullptr, 0) * 2654435761u + 12345u;
    if (!p.w || !p.h) { err = "bad size in " + spec; return false; }
    const std::string& kind = f[1];
    const size_t n = (size_t)p.w * p.h;
    if (kind == "sky") {
        p.is_float = true;
        p.ch = 4;
        p.f32.resize(n * 4);
        for (uint32_t y = 0; y < p.h; y++)
            for (uint32_t x = 0; x < p.w; x++) {
                const float v = (float)y / p.h, u = (float)x / p.w;
                const float sun = expf(-60.0f * ((u - 0.7f) * (u - 0.7f) + (v - 0.3f) * (v - 0.3f))) * 40.0f;
                float* o = &p.f32[((size_t)y * p.w + x) * 4];
                o[0] = 0.2f + 0.5f * v + sun + (lcg(seed) & 255) / 4096.0f;
                o[1] = 0.3f + 0.4f * v + sun * 0.9f;
                o[2] = 0.9f - 0.5f * v + sun * 0.6f;
                o[3] = 1.0f;
            }
        return true;
    }
    p.ch = kind == "rough" ? 1 : kind == "normal" ? 2 : 4;
    if (kind != "rough" && kind != "normal" && kind != "albedo" && kind != "mask") { err = "unknown kind '" + kind + "'"; return false; }
    p.u8.resize(n * (size_t)p.ch);
    for (uint32_t y = 0; y < p.h; y++)
        for (uint32_t x = 0; x < p.w; x++) {
            uint8_t* o = &p.u8[((size_t)y * p.w + x) * (size_t)p.ch];
            const float u = (float)x / p.w, v = (float)y / p.h;
            const int nz = (int)(lcg(seed) & 15);
            if (p.ch == 1) o[0] = (uint8_t)(128 + 100 * sinf(20 * u) * cosf(14 * v) + nz);
            else if (p.ch == 2) { o[0] = (uint8_t)(128 + 90 * sinf(31 * u + 3 * v)); o[1] = (uint8_t)(128 + 90 * cosf(27 * v - 2 * u)); }
            else {
                o[0] = (uint8_t)(40 + 180 * u + nz);
                o[1] = (uint8_t)(60 + 150 * v + nz);
                o[2] = (uint8_t)(128 + 100 * sinf(25 * (u + v)));
                o[3] = kind == "mask" ? (uint8_t)((((x >> 4) ^ (y >> 4)) & 1) ? 255 : (16 * ((x + y) & 15))) : 255;
            }
        }
    return true;
}

bool read_pnm(const std::string& path, Pixels& p, std::string& err)
{
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) { err = "Failed to load " + path; return false; }
    char magic[3] = {0, 0, 0};
    auto token = [&](std::string& out) {
        out.clear();
        int c;
        for (;;) {
            c = fgetc(f);
            if (c == '#') { while (c != '\n' && c != EOF) c = fgetc(f); continue; }
            if (c == EOF || !isspace(c)) break;
        }
        while (c != EOF && !isspace(c)) { out.push_back((char)c); c = fgetc(f); }
        return !out.empty();
    };
    std::string t;
    bool ok = fread(magic, 1, 2, f) == 2;
    const std::string m(magic);
    if (ok && (m == "P5" || m == "P6")) {
        std::string a, b, c;
        ok = token(a) && token(b) && token(c);
        p.w = (uint32_t)atoi(a.c_str());
        p.h = (uint32_t)atoi(b.c_str());
        p.ch = m == "P5" ? 1 : 3;
        ok = ok && atoi(c.c_str()) == 255 && p.w && p.h;
        if (ok) {
            p.u8.resize((size_t)p.w * p.h * (size_t)p.ch);
            ok = fread(p.u8.data(), 1, p.u8.size(), f) == p.u8.size();
        }
    } else if (ok && (m == "Pf" || m == "PF")) {
        std::string a, b, c;
        ok = token(a) && token(b) && token(c);
        p.w = (uint32_t)atoi(a.c_str());
        p.h = (uint32_t)atoi(b.c_str());
        p.ch = m == "Pf" ? 1 : 3;
        p.is_float = true;
        const double scale = atof(c.c_str());
        ok = ok && scale < 0 && p.w && p.h; // little-endian files only
        if (ok) {
            std::vector<float> rows((size_t)p.w * p.h * (size_t)p.ch);
            ok = fread(rows.data(), 4, rows.size(), f) == rows.size();
            p.f32.resize(rows.size());
            const size_t rb = (size_t)p.w * (size_t)p.ch;
            for (uint32_t y = 0; y < p.h && ok; y++) // PFM stores the bottom row first
                memcpy(&p.f32[(size_t)y * rb], &rows[(size_t)(p.h - 1 - y) * rb], rb * 4);
        }
    } else ok = false;
    fclose(f);
    if (!ok) err = "Failed to load " + path + " (binary P5 / P6 with maxval 255, or little-endian Pf / PF expected)";
    return ok;
}

// source texels for `fmt` (hkbcn_src_texel_bytes): RGBA8 / R8 / RG8 / RGBA32F
// (pinned host memory: the batch then uploads straight from it with the copy engine)
bool to_source(const Pixels& p, int fmt, uint8_t*& out, std::string& err)
{
    const size_t n = (size_t)p.w * p.h;
    const int bpp = hkbcn_src_texel_bytes(fmt, 0);
    out = (uint8_t*)hkbcn_alloc_pinned(n * (size_t)bpp);
    if (!out) { err = hkbcn_last_error(); return false; }
    if (bpp == 16) {
        if (!p.is_float) { err = "BC6H needs a float source (PFM or gen:sky)"; return false; }
        float* o = (float*)out;
        for (size_t i = 0; i < n; i++)
            for (int c = 0; c < 4; c++) o[i * 4 + c] = c < p.ch ? p.f32[i * (size_t)p.ch + c] : (c == 3 ? 1.0f : (p.ch == 1 ? p.f32[i] : 0.0f));
        return true;
    }
    if (p.is_float) { err = "an 8-bit format needs an 8-bit source"; return false; }
    const int oc = bpp; // 1, 2 or 4 channels of one byte
    for (size_t i = 0; i < n; i++)
        for (int c = 0; c < oc; c++) {
            uint8_t v;
            if (c < p.ch) v = p.u8[i * (size_t)p.ch + c];
            else if (c == 3) v = 255;
            else v = p.ch == 1 ? p.u8[i] : 0; // grey -> RGB
            out[i * (size_t)oc + c] = v;
        }
    return true;
}

std::vector<std::string> split_ws(const std::string& line)
{
    std::vector<std::string> t;
    std::stringstream ss(line);
    for (std::string x; ss >> x;) t.push_back(x);
    return t;
}

// ---- CubemapImporter (Tools/CubemapImporter/Main.cpp:58-145 -> LoadSkyboxImages, Image.cpp:1918-2040) -------------------
//     -cubemap -s <six face files: +X -X +Y -Y +Z -Z> -o <file> -format BC6H_UFLOAT | BC1_UNORM_SRGB [-hdri_scale v] [-hdri_pow v]
// The six faces are compressed in ONE launch (hkbcn_compress_slices, ImageStorage order) and written as a TEXTURE_CUBE
// storage of six slices, one mip, IMAGE_STORAGE_NO_ALPHA (Image.cpp:2002-2011).  The reference's uncompressed skybox
// formats (SRGBA8 / SBGRA8 / R11G11B10F) are not block compression and stay out; faces whose size is not a multiple of 4
// would need its Mitchell / clamp pre-resample (Image.cpp:1960-1982), which is the CPU -resample pre-pass left out above.
int run_cubemap(const std::vector<std::string>& t, uint32_t mask)
{
    std::string faces[6], out, format = "BC6H_UFLOAT";
    float scale = 1.0f, pw = 1.0f;
    int nf = 0;
    for (size_t i = 0; i < t.size(); i++) {
        if (t[i] == "-cubemap") continue;
        if (t[i] == "-s" && i + 6 < t.size()) { for (nf = 0; nf < 6; nf++) faces[nf] = t[++i]; }
        else if (t[i] == "-o" && i + 1 < t.size()) out = t[++i];
        else if (t[i] == "-format" && i + 1 < t.size()) format = t[++i];
        else if (t[i] == "-hdri_scale" && i + 1 < t.size()) scale = (float)atof(t[++i].c_str());
        else if (t[i] == "-hdri_pow" && i + 1 < t.size()) pw = (float)atof(t[++i].c_str());
        else { fprintf(stderr, "unknown or incomplete flag '%s'\n", t[i].c_str()); return -1; }
    }
    if (out.empty()) { fprintf(stderr, "Output file is not specified. Use -o <filename>\n"); return -1; }
    if (nf != 6) { fprintf(stderr, "Source files not specified. Use -s <cubemap faces>\n"); return -1; }
    const int fmt = find_format(format);
    if (fmt != HKBCN_BC6H_UFLOAT && fmt != HKBCN_BC1_UNORM_SRGB) { fprintf(stderr, "Unexpected texture format %s\n", format.c_str()); return -1; }
    if (hkbcn_init(mask) != HKBCN_OK) { fprintf(stderr, "%s\n", hkbcn_last_error()); return -1; }
    const bool hdri = fmt == HKBCN_BC6H_UFLOAT;
    uint32_t w = 0;
    uint8_t* packed = nullptr;
    size_t face_bytes = 0;
    for (int f = 0; f < 6; f++) {
        Pixels p;
        std::string err;
        printf("Loading %s...\n", faces[f].c_str());
        const bool ok = faces[f].rfind("gen:", 0) == 0 ? generate(faces[f], p, err) : read_pnm(faces[f], p, err);
        if (!ok) { fprintf(stderr, "%s\n", err.c_str()); return -1; }
        if (f == 0) w = p.w;
        if (p.w != w || p.w != p.h) { fprintf(stderr, "LoadSkyboxImages: Invalid image size\n"); return -1; }
        if (w & 3) { fprintf(stderr, "face size %u is not a multiple of the block size 4 (pre-resample is not part of this tool)\n", w); return -1; }
        if (hdri && p.is_float && (scale != 1.0f || pw != 1.0f)) // Image.cpp:1984-2000
            for (size_t j = 0; j < p.f32.size(); j += (size_t)p.ch)
                for (int c = 0; c < 3 && c < p.ch; c++) p.f32[j + (size_t)c] = powf(p.f32[j + (size_t)c] * scale, pw);
        uint8_t* one = nullptr;
        if (!to_source(p, fmt, one, err)) { fprintf(stderr, "%s\n", err.c_str()); return -1; }
        if (f == 0) {
            face_bytes = (size_t)w * w * (size_t)hkbcn_src_texel_bytes(fmt, 0);
            packed = (uint8_t*)hkbcn_alloc_pinned(6 * face_bytes);
            if (!packed) { fprintf(stderr, "%s\n", hkbcn_last_error()); return -1; }
        }
        memcpy(packed + (size_t)f * face_bytes, one, face_bytes);
        hkbcn_free_pinned(one);
    }
    // one cube job through the batch entry point: six slices, one mip, IMAGE_STORAGE_NO_ALPHA; the library compresses the six
    // faces in one launch and writes the TEXTURE_CUBE storage
    printf("Importing texture %s...\n", out.c_str());
    const auto t0 = std::chrono::steady_clock::now();
    hkbcn_import_job job;
    memset(&job, 0, sizeof(job));
    job.fmt = fmt;
    job.width = job.height = w;
    job.slices = 6;
    job.mip.no_alpha = 1;
    job.src_level0 = packed;
    job.out_path = out.c_str();
    if (hkbcn_import_batch(&job, 1) != HKBCN_OK) { fprintf(stderr, "import failed: %s\n", hkbcn_last_error()); return -1; }
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    printf("cubemap 6 x %u^2, %.1f Mpix in %.3f s: %.2f Mpix/s including upload, download and the file write\n", w, 6.0 * w * w / 1e6, s,
           6.0 * w * w / 1e6 / s);
    hkbcn_free_pinned(packed);
    hkbcn_shutdown();
    return 0;
}

// ---- PBRImporter (Tools/PBRImporter/Main.cpp:284-755) -----------------------------------------------------------------
//     -pbr_preset ORMX_BC1|ORMX_BC3|ORMX_BC7 <out> [-occlusion f ch] [-roughness f ch] [-metallic f ch]
//     [-displacement_to_alpha f ch | -emission_to_alpha f ch] [-normal <src> <dst>] [-a vMF|Toksvig] [-dx_normal]
//     [-normal_pack <type>] [-mip_edge_mode m] [-mip_filter f]
// ORMX: the four maps are gathered into one RGBA8 image (CreateORMX, Main.cpp:86-190: defaults 255 255 0 255), its mips are
// generated as an RGBA8_UNORM storage with IMAGE_STORAGE_NO_ALPHA (:596-613), the roughness channel of every level >= 1 is
// updated from the triangle-filtered normals (:616-680), the chain is compressed and written (:683-735).  Normal map:
// v*2-1 + normalise (:197-206), CreateNormalMap (Image.cpp:2262-2346): pack level 0, mip the PACKED texels (triangle filter,
// NO_ALPHA), compress every level.  Every step is a libhkbcn call; what stays out is what stays on the CPU in the reference:
// file decoding, -resample, and the uncompressed presets (ORMX, -no_compress_normals), which are not block compression.
struct MapArg { std::string file; int ch = 0; };

bool load_rgba8(const std::string& src, Pixels& p, std::vector<uint8_t>& rgba, std::string& err)
{
    if (!(src.rfind("gen:", 0) == 0 ? generate(src, p, err) : read_pnm(src, p, err))) return false;
    if (p.is_float) { err = src + ": an 8-bit source is expected"; return false; }
    const size_t n = (size_t)p.w * p.h;
    rgba.resize(n * 4);
    for (size_t i = 0; i < n; i++)
        for (int c = 0; c < 4; c++) rgba[i * 4 + (size_t)c] = c < p.ch ? p.u8[i * (size_t)p.ch + (size_t)c] : (c == 3 ? 255 : (p.ch == 1 ? p.u8[i] : 0));
    return true;
}

int run_pbr(const std::vector<std::string>& t, uint32_t mask)
{
    MapArg maps[4]; // occlusion, roughness, metallic, displacement / emission
    maps[0].ch = 1; maps[1].ch = 1; maps[2].ch = 0; maps[3].ch = 0; // the reference's default channel indices
    std::string preset, out_pbr, normal_src, normal_dst;
    int bake = 0, pack = HKBCN_NORMALS_RGBA_BC1, edge = 0, filter = 0;
    bool mips = false, dx = false;
    for (size_t i = 0; i < t.size(); i++) {
        auto need = [&](int n) { return i + (size_t)n < t.size(); };
        int which = t[i] == "-occlusion" ? 0 : t[i] == "-roughness" ? 1 : t[i] == "-metallic" ? 2 : (t[i] == "-displacement_to_alpha" || t[i] == "-emission_to_alpha") ? 3 : -1;
        if (which >= 0 && need(2)) {
            maps[which].file = t[++i];
            maps[which].ch = atoi(t[++i].c_str());
            if (maps[which].ch < 0 || maps[which].ch >= 4) { fprintf(stderr, "Invalid channel index %d\n", maps[which].ch); return -1; }
        } else if (t[i] == "-pbr_preset" && need(2)) { preset = t[++i]; out_pbr = t[++i]; }
        else if (t[i] == "-normal" && need(2)) { normal_src = t[++i]; normal_dst = t[++i]; }
        else if (t[i] == "-a" && need(1)) { ++i; bake = (t[i] == "vMF" || t[i] == "vmf") ? HKBCN_ROUGHNESS_BAKE_VMF : (t[i] == "Toksvig" || t[i] == "toksvig") ? HKBCN_ROUGHNESS_BAKE_TOKSVIG : 0; }
        else if (t[i] == "-dx_normal") dx = true;
        else if (t[i] == "-normal_pack" && need(1)) {
            static const char* names[6] = {"RGBA_BC1_COMPATIBLE", "RG_BC5_COMPATIBLE", "SPHEREMAP_BC5_COMPATIBLE", "STEREOGRAPHIC_BC5_COMPATIBLE",
                                           "PARABOLOID_BC5_COMPATIBLE", "RGBA_BC3_COMPATIBLE"};
            ++i;
            pack = HKBCN_NORMALS_RGBA_BC1; // GetNormalMapPack falls back to RGBA_BC1_COMPATIBLE on an unknown name
            for (int k = 0; k < 6; k++)
                if (t[i] == names[k]) pack = k;
        } else if (t[i] == "-mip_edge_mode" && need(1)) { edge = edge_mode(t[++i]); mips = true; }
        else if (t[i] == "-mip_filter" && need(1)) { filter = filter_of(t[++i]); mips = true; }
        else { fprintf(stderr, "unknown or incomplete flag '%s'\n", t[i].c_str()); return -1; }
    }
    int fmt = -1;
    if (!preset.empty()) {
        fmt = preset == "ORMX_BC1" ? HKBCN_BC1_UNORM : preset == "ORMX_BC3" ? HKBCN_BC3_UNORM : preset == "ORMX_BC7" ? HKBCN_BC7_UNORM : -1;
        if (fmt < 0) { fprintf(stderr, "Unknown PBR preset: %s (the uncompressed ORMX preset is not block compression)\n", preset.c_str()); return -1; }
    }
    if (preset.empty() && normal_src.empty()) { fputs("nothing to do: give -pbr_preset and / or -normal\n", stderr); return -1; }
    if (hkbcn_init(mask) != HKBCN_OK) { fprintf(stderr, "%s\n", hkbcn_last_error()); return -1; }
    auto check_dims = [&](uint32_t w, uint32_t h, const char* what) {
        if ((w & 3) || (h & 3) || (mips && ((w & (w - 1)) || (h & (h - 1))))) {
            fprintf(stderr, "%s: %u x %u must be a multiple of 4%s (resampling is not part of this tool)\n", what, w, h, mips ? " and a power of two" : "");
            return false;
        }
        return true;
    };
    std::string err;
    // ---- normal map source: Float3 in [0, 1] -> v * 2 - 1, normalised
    std::vector<float> normals;
    uint32_t nw = 0, nh = 0;
    if (!normal_src.empty()) {
        Pixels p;
        printf("Loading %s...\n", normal_src.c_str());
        if (!(normal_src.rfind("gen:", 0) == 0 ? generate(normal_src, p, err) : read_pnm(normal_src, p, err))) { fprintf(stderr, "%s\n", err.c_str()); return -1; }
        if (p.ch < 2) { fprintf(stderr, "%s: a normal map needs at least two channels\n", normal_src.c_str()); return -1; }
        nw = p.w; nh = p.h;
        if (!check_dims(nw, nh, normal_src.c_str())) return -1;
        const size_t n = (size_t)nw * nh;
        normals.resize(n * 3);
        for (size_t i = 0; i < n; i++)
            for (int c = 0; c < 3; c++) {
                float v = c < p.ch ? (p.is_float ? p.f32[i * (size_t)p.ch + (size_t)c] : p.u8[i * (size_t)p.ch + (size_t)c] / 255.0f) : 1.0f;
                if (dx && c == 1) v = 1.0f - v; // RawImage::InvertGreen
                normals[i * 3 + (size_t)c] = v;
            }
        if (hkbcn_normalize_vectors(normals.data(), n) != HKBCN_OK) { fprintf(stderr, "%s\n", hkbcn_last_error()); return -1; }
    }
    const auto t0 = std::chrono::steady_clock::now();
    double mpix = 0;
    // ---- ORMX
    if (fmt >= 0) {
        // one hkbcn_pbr_job: gather, cascade, roughness update, encode and the file all happen behind hkbcn_import_pbr_batch
        std::vector<uint8_t> rgba[4];
        uint32_t w = 0, h = 0;
        bool any = false;
        hkbcn_pbr_job job;
        memset(&job, 0, sizeof(job));
        for (int k = 0; k < 4; k++) {
            job.channel[k] = (uint8_t)maps[k].ch;
            if (maps[k].file.empty()) continue;
            Pixels p;
            printf("Loading %s...\n", maps[k].file.c_str());
            if (!load_rgba8(maps[k].file, p, rgba[k], err)) { fprintf(stderr, "%s\n", err.c_str()); return -1; }
            if (!any) { w = p.w; h = p.h; any = true; }
            else if (p.w != w || p.h != h) {
                fputs("The source images have different sizes. Use the -resample option to configure resampling parameters.\n", stderr);
                return -1;
            }
            job.maps[k] = rgba[k].data();
        }
        if (!any) { fputs("-pbr_preset needs at least one source map\n", stderr); return -1; }
        if (!check_dims(w, h, "ORMX")) return -1;
        if (bake && !normals.empty() && (w != nw || h != nh)) { fputs("Can't apply antialiasing. ORMX and NormalMap must be the same size.\n", stderr); return -1; }
        job.fmt = fmt;
        job.width = w; job.height = h;
        job.normals = (bake && !normals.empty()) ? normals.data() : nullptr;
        job.roughness_bake = job.normals ? bake : 0;
        job.generate_mips = mips ? 1 : 0;
        job.mip.edge_mode = edge; job.mip.filter = filter;
        job.out_path = out_pbr.c_str();
        printf("Importing texture %s...\n", out_pbr.c_str());
        if (hkbcn_import_pbr_batch(&job, 1) != HKBCN_OK) { fprintf(stderr, "%s\n", hkbcn_last_error()); return -1; }
        mpix += (double)w * h * (mips ? 4.0 / 3.0 : 1.0) / 1e6;
    }
    // ---- normal map
    if (!normals.empty()) {
        const int nfmt = pack == HKBCN_NORMALS_RGBA_BC1 ? HKBCN_BC1_UNORM : pack == HKBCN_NORMALS_RGBA_BC3 ? HKBCN_BC3_UNORM : HKBCN_BC5_UNORM;
        const size_t n = (size_t)nw * nh;
        std::vector<uint8_t> packed(n * (size_t)hkbcn_pack_texel_bytes(pack));
        if (hkbcn_pack_normals(pack, normals.data(), packed.data(), n) != HKBCN_OK) { fprintf(stderr, "%s\n", hkbcn_last_error()); return -1; }
        const uint32_t nm = mips ? hkbcn_calc_num_mips(nw, nh) : 1;
        hkbcn_mip_config cfg = {edge, 2 /* IMAGE_RESAMPLE_FILTER_TRIANGLE, Main.cpp:748 */, 1, 0};
        std::vector<uint8_t> blocks(hkbcn_chain_bytes(nfmt, nw, nh, 1, nm));
        printf("Importing texture %s...\n", normal_dst.c_str());
        if (hkbcn_import_chain(nfmt, packed.data(), blocks.data(), nw, nh, nm, &cfg, 0) != HKBCN_OK) { fprintf(stderr, "%s\n", hkbcn_last_error()); return -1; }
        hkbcn_texture_desc d = {HKBCN_TEXTURE_2D, nw, nh, 1, nm, nfmt, HKBCN_STORAGE_NO_ALPHA};
        if (hkbcn_write_texture(normal_dst.c_str(), &d, blocks.data(), blocks.size()) != HKBCN_OK) { fprintf(stderr, "%s\n", hkbcn_last_error()); return -1; }
        mpix += (double)n * (mips ? 4.0 / 3.0 : 1.0) / 1e6;
    }
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    printf("PBR import: %.1f Mpix in %.3f s (%.1f Mpix/s including transfers and file writes)\n", mpix, sec, mpix / sec);
    hkbcn_shutdown();
    return 0;
}

const char* kHelp = R"(hk_texture_import -- BCn texture import on the GPU (libhkbcn)
    -h                      -- Help
    -s <filename>           -- Source: binary PGM/PPM (P5/P6), PFM (Pf/PF) or gen:<albedo|mask|rough|normal|sky>:<w>:<h>:<seed>
    -o <filename>           -- Output .texture file
    -no_alpha               -- Don't aware about alpha channel or ignore it
    -alpha_premult          -- Set this flag if your texture has premultiplied alpha
    -format <name>          -- Output texture format: BC1_UNORM[_SRGB] BC2_.. BC3_.. BC4_UNORM BC5_UNORM BC6H_UFLOAT BC6H_SFLOAT BC7_UNORM[_SRGB]
    -mip_edge_mode <mode>   -- clamp | reflect | wrap | zero          (giving either -mip_* flag generates the mip chain)
    -mip_filter <filter>    -- box | triangle | cubicspline | catmullrom | mitchell
    -resample <width> <height>                  -- Scale input textures (on the device; rounded to a power of two with mips, else to 4)
    -resample_edge_mode <mode_h> <mode_v>       -- Use edge mode for resampling (clamp/reflect/wrap/zero)
    -resample_filter <filter_h> <filter_v>      -- Use filter for resampling (box/triangle/cubicspline/catmullrom/mitchell)
    -batch <manifest>       -- one flag set per line; all lines are imported as one batch sharded over the GPUs
    -gpus <mask>            -- bit mask of the CUDA devices to use (default: all)
    -skip_existing          -- leave textures alone whose output file is already complete (re-run of an interrupted batch)
    -pbr_preset / -normal   -- PBRImporter mode: -pbr_preset ORMX_BC1|ORMX_BC3|ORMX_BC7 <out> [-occlusion f ch] [-roughness f ch] [-metallic f ch]
                               [-displacement_to_alpha f ch | -emission_to_alpha f ch] [-normal <src> <dst>] [-a vMF|Toksvig] [-dx_normal]
                               [-normal_pack <type>] [-mip_edge_mode m] [-mip_filter f]
    -cubemap                -- CubemapImporter mode: -s <six faces +X -X +Y -Y +Z -Z> -o <file> -format BC6H_UFLOAT | BC1_UNORM_SRGB
                               [-hdri_scale <v>] [-hdri_pow <v>]   (result = pow(color * hdri_scale, hdri_pow))
)";

} // namespace

int main(int argc, char** argv)
{
    std::vector<std::string> argvv(argv + 1, argv + argc);
    std::vector<Args> all;
    uint32_t mask = 0;
    bool skip_existing = false;
    std::vector<std::string> single;
    for (size_t i = 0; i < argvv.size(); i++) {
        if (argvv[i] == "-h") { fputs(kHelp, stdout); return 0; }
        if (argvv[i] == "-skip_existing") { skip_existing = true; continue; }
        if (argvv[i] == "-gpus" && i + 1 < argvv.size()) { mask = (uint32_t)strtoul(argvv[++i].c_str(), nullptr, 0); continue; }
        if (argvv[i] == "-pbr_preset" || argvv[i] == "-normal") {
            std::vector<std::string> rest;
            for (size_t k = 0; k < argvv.size(); k++) {
                if (argvv[k] == "-gpus" && k + 1 < argvv.size()) { mask = (uint32_t)strtoul(argvv[++k].c_str(), nullptr, 0); continue; }
                rest.push_back(argvv[k]);
            }
            return run_pbr(rest, mask);
        }
        if (argvv[i] == "-cubemap") {
            std::vector<std::string> rest;
            for (size_t k = 0; k < argvv.size(); k++) {
                if (argvv[k] == "-gpus" && k + 1 < argvv.size()) { mask = (uint32_t)strtoul(argvv[++k].c_str(), nullptr, 0); continue; }
                rest.push_back(argvv[k]);
            }
            return run_cubemap(rest, mask);
        }
        if (argvv[i] == "-batch" && i + 1 < argvv.size()) {
            std::ifstream mf(argvv[++i]);
            if (!mf) { fprintf(stderr, "Failed to open manifest %s\n", argvv[i].c_str()); return -1; }
            int ln = 0;
            for (std::string line; std::getline(mf, line);) {
                ln++;
                const size_t hash = line.find('#');
                if (hash != std::string::npos) line.resize(hash);
                const auto t = split_ws(line);
                if (t.empty()) continue;
                Args a;
                std::string err;
                if (!parse_flags(t, a, err)) { fprintf(stderr, "%s:%d: %s\n", argvv[i].c_str(), ln, err.c_str()); return -1; }
                all.push_back(a);
            }
            continue;
        }
        single.push_back(argvv[i]);
    }
    if (!single.empty()) {
        Args a;
        std::string err;
        if (!parse_flags(single, a, err)) { fprintf(stderr, "%s\n", err.c_str()); return -1; }
        all.push_back(a);
    }
    if (all.empty()) { fputs(kHelp, stdout); return -1; }
    if (hkbcn_init(mask) != HKBCN_OK) { fprintf(stderr, "%s\n", hkbcn_last_error()); return -1; }

    std::vector<uint8_t*> sources(all.size(), nullptr);
    std::vector<hkbcn_import_job> jobs(all.size());
    double mpix = 0;
    for (size_t i = 0; i < all.size(); i++) {
        const Args& a = all[i];
        Pixels p;
        std::string err;
        printf("Loading %s...\n", a.src.c_str());
        const bool ok = a.src.rfind("gen:", 0) == 0 ? generate(a.src, p, err) : read_pnm(a.src, p, err);
        const int fmt = find_format(a.format);
        if (!ok || !to_source(p, fmt, sources[i], err)) { fprintf(stderr, "%s\n", err.c_str()); return -1; }
        if (a.resample) {
            // Main.cpp:163-178: the requested size goes to the closest power of two (at least one block) when mips are generated,
            // up to the next multiple of the block size otherwise; alpha weighting and colour space follow the format
            auto pow2 = [](uint32_t v) { uint32_t hi = 1; while (hi < v) hi <<= 1; const uint32_t lo = hi >> 1; return (hi - v) < (v - lo) ? hi : lo; };
            hkbcn_resample_params rp;
            memset(&rp, 0, sizeof(rp));
            rp.scaled_width = a.mips ? std::max<uint32_t>(pow2(a.rs_w), 4) : (a.rs_w + 3) & ~3u;
            rp.scaled_height = a.mips ? std::max<uint32_t>(pow2(a.rs_h), 4) : (a.rs_h + 3) & ~3u;
            rp.edge_h = a.rs_edge_h; rp.edge_v = a.rs_edge_v; rp.filter_h = a.rs_filter_h; rp.filter_v = a.rs_filter_v;
            const int bpp = hkbcn_src_texel_bytes(fmt, 0);
            const bool has_alpha = bpp == 4; // TexFormat[].bHasAlpha: BC1 / 2 / 3 / 7
            const bool srgb = fmt == HKBCN_BC1_UNORM_SRGB || fmt == HKBCN_BC2_UNORM_SRGB || fmt == HKBCN_BC3_UNORM_SRGB || fmt == HKBCN_BC7_UNORM_SRGB;
            if (has_alpha && !a.no_alpha) rp.flags |= HKBCN_RESAMPLE_HAS_ALPHA | (a.alpha_premult ? HKBCN_RESAMPLE_ALPHA_PREMULTIPLIED : 0);
            if (srgb) rp.flags |= HKBCN_RESAMPLE_COLORSPACE_SRGB;
            const int raw = bpp == 1 ? HKBCN_RAW_R8 : bpp == 2 ? HKBCN_RAW_R8_ALPHA : bpp == 4 ? HKBCN_RAW_RGBA8 : HKBCN_RAW_RGBA32F;
            uint8_t* scaled = (uint8_t*)hkbcn_alloc_pinned((size_t)rp.scaled_width * rp.scaled_height * (size_t)bpp);
            if (!scaled || hkbcn_resample_raw_image(raw, sources[i], p.w, p.h, &rp, scaled) != HKBCN_OK) { fprintf(stderr, "%s\n", hkbcn_last_error()); return -1; }
            hkbcn_free_pinned(sources[i]);
            sources[i] = scaled;
            p.w = rp.scaled_width;
            p.h = rp.scaled_height;
        }
        // CalcNumMips / ImageStorage (Image.cpp:203-204, :415-416)
        if ((p.w & 3) || (p.h & 3)) { fprintf(stderr, "%s: width and height must be multiples of the block size 4\n", a.src.c_str()); return -1; }
        if (a.mips && ((p.w & (p.w - 1)) || (p.h & (p.h - 1)))) {
            fprintf(stderr, "%s: width and height must be powers of two for compressed mipmapped textures\n", a.src.c_str());
            return -1;
        }
        hkbcn_import_job& j = jobs[i];
        memset(&j, 0, sizeof(j));
        j.fmt = fmt;
        j.width = p.w;
        j.height = p.h;
        j.generate_mips = a.mips ? 1 : 0;
        j.mip.edge_mode = a.edge;
        j.mip.filter = a.filter;
        j.mip.no_alpha = a.no_alpha ? 1 : 0;
        j.mip.alpha_premultiplied = a.alpha_premult ? 1 : 0;
        j.src_level0 = sources[i];
        j.out_path = a.out.c_str();
        if (skip_existing) j.flags |= HKBCN_FLAG_SKIP_EXISTING;
        mpix += (double)p.w * p.h * (a.mips ? 4.0 / 3.0 : 1.0) / 1e6;
    }
    for (auto& a : all) printf("Importing texture %s...\n", a.out.c_str());
    const auto t0 = std::chrono::steady_clock::now();
    const int rc = hkbcn_import_batch(jobs.data(), jobs.size());
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (rc != HKBCN_OK) { fprintf(stderr, "import failed: %s\n", hkbcn_last_error()); return -1; }
    size_t skipped = 0;
    for (auto& j : jobs) skipped += j.status == HKBCN_SKIPPED;
    if (skipped) printf("%zu texture(s) already complete, skipped\n", skipped);
    printf("%zu texture(s), %.1f Mpix in %.3f s on %d GPU(s): %.1f Mpix/s including upload, download and file writes\n", jobs.size(), mpix,
           s, hkbcn_device_count(), mpix / s);
    for (uint8_t* q : sources) hkbcn_free_pinned(q);
    hkbcn_shutdown();
    return 0;
}
