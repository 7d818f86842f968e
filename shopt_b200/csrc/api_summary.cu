// api_summary.cu -- summary.shopt, the on-disk product of the stage.
//
// Replaces the astropy calls of shopt/dataOutprocessing.jl:52-84 (truncate_summary_file: true, the shipped default) and
// shopt/reader.jl:13-38 (read_shopt) with a self-contained FITS writer / reader (host code):
//   HDU 0  primary image  POLYNOMIAL_MATRIX   float64, numpy shape (r, c, K)  ->  NAXIS1 = K, NAXIS2 = c, NAXIS3 = r
//   HDU 1  BINTABLE       SUMMARY_STATISTICS  column POLYNOMIAL_DEGREE (D), one row
//   HDU 2  BINTABLE       ANALYTIC_PROFILE    columns s_MATRIX, g1_MATRIX, g2_MATRIX (D), ten rows: the cubic field fits
// shopt_summary_write_full writes the seven-HDU layout of dataOutprocessing.jl:86-168 (truncate_summary_file: false):
//   0 POLYNOMIAL_MATRIX | 1 SUMMARY_STATISTICS (13 columns, zero-padded to the longest, like astropy's from_columns)
//   2 VIGNETS, numpy (r, c, N) -- the catalogue was stacked into array_3d first (:29-34)
//   3 PIXEL_GRID_FITS, numpy (N, r, c) -- the py block interpolates the Vector{Matrix} `pixelGridFits`, not the stacked
//     `pixelgridfits` of :36-40 (different spelling), so numpy stacks the matrices along a NEW FIRST axis
//   4 VALIDATION_STARS, numpy (N_valid, r, c), same reason | 5 FLAGS (3 columns, zero-padded) | 6 ANALYTIC_PROFILE
// FITS: 2880-byte blocks, 80-character header cards, big-endian data.  The reader locates columns by NAME in whatever
// HDU carries them, so the reference's full (non-truncated) layout -- ANALYTIC_PROFILE is HDU 6 there -- reads too.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "api_ctx.h"

namespace shopt {

static void put_card(std::string &h, const char *key, const std::string &value, bool quoted = false) {
    char card[81];
    if (quoted) snprintf(card, sizeof card, "%-8s= '%-8s'", key, value.c_str());
    else snprintf(card, sizeof card, "%-8s= %20s", key, value.c_str());
    std::string c(card);
    c.resize(80, ' ');
    h += c;
}
static void end_header(std::string &h) {
    std::string e = "END";
    e.resize(80, ' ');
    h += e;
    while (h.size() % 2880) h += ' ';
}
static void put_be64(std::string &d, double x) {
    unsigned long long b;
    memcpy(&b, &x, 8);
    for (int s = 56; s >= 0; s -= 8) d += (char)((b >> s) & 0xff);
}
static double get_be64(const unsigned char *p) {
    unsigned long long b = 0;
    for (int i = 0; i < 8; i++) b = (b << 8) | p[i];
    double x;
    memcpy(&x, &b, 8);
    return x;
}
static void pad_data(std::string &d) { while (d.size() % 2880) d += '\0'; }

static std::string bintable(const char *extname, const std::vector<std::pair<std::string, const double *>> &cols, long long rows) {
    std::string h, d;
    put_card(h, "XTENSION", "BINTABLE", true);
    put_card(h, "BITPIX", "8");
    put_card(h, "NAXIS", "2");
    put_card(h, "NAXIS1", std::to_string(8 * cols.size()));
    put_card(h, "NAXIS2", std::to_string(rows));
    put_card(h, "PCOUNT", "0");
    put_card(h, "GCOUNT", "1");
    put_card(h, "TFIELDS", std::to_string(cols.size()));
    for (size_t i = 0; i < cols.size(); i++) {
        put_card(h, ("TTYPE" + std::to_string(i + 1)).c_str(), cols[i].first, true);
        put_card(h, ("TFORM" + std::to_string(i + 1)).c_str(), "D", true);
    }
    put_card(h, "EXTNAME", extname, true);
    end_header(h);
    for (long long r = 0; r < rows; r++)
        for (auto &c : cols) put_be64(d, c.second[r]);
    pad_data(d);
    return h + d;
}

// ---- reading ------------------------------------------------------------------------------------------------
struct Hdu {
    std::vector<std::pair<std::string, std::string>> cards;   // key, raw value
    size_t data_off = 0, data_len = 0;
    std::string get(const std::string &k) const {
        for (auto &c : cards) if (c.first == k) return c.second;
        return "";
    }
    long long geti(const std::string &k, long long def = 0) const {
        const std::string v = get(k);
        return v.empty() ? def : atoll(v.c_str());
    }
};
static std::string trim_value(std::string v) {
    const size_t q = v.find('\'');
    if (q != std::string::npos) {                              // quoted string
        const size_t e = v.find('\'', q + 1);
        v = v.substr(q + 1, e == std::string::npos ? std::string::npos : e - q - 1);
        while (!v.empty() && v.back() == ' ') v.pop_back();
        return v;
    }
    const size_t sl = v.find('/');
    if (sl != std::string::npos) v = v.substr(0, sl);
    size_t a = 0, b = v.size();
    while (a < b && v[a] == ' ') a++;
    while (b > a && v[b - 1] == ' ') b--;
    return v.substr(a, b - a);
}
static bool parse_hdus(const std::vector<unsigned char> &f, std::vector<Hdu> &out) {
    size_t off = 0;
    while (off + 2880 <= f.size()) {
        Hdu h;
        bool ended = false;
        while (!ended) {
            if (off + 2880 > f.size()) return false;
            for (int c = 0; c < 36; c++) {
                std::string card((const char *)&f[off + 80 * c], 80);
                std::string key = card.substr(0, 8);
                while (!key.empty() && key.back() == ' ') key.pop_back();
                if (key == "END") { ended = true; break; }
                if (card[8] == '=') h.cards.push_back({key, trim_value(card.substr(10))});
            }
            off += 2880;
        }
        const long long naxis = h.geti("NAXIS");
        long long n = naxis > 0 ? 1 : 0;
        for (long long a = 1; a <= naxis; a++) n *= h.geti("NAXIS" + std::to_string(a));
        const long long bitpix = h.geti("BITPIX");
        long long bytes = (bitpix < 0 ? -bitpix : bitpix) / 8 * h.geti("GCOUNT", 1) * (h.geti("PCOUNT", 0) + n);
        h.data_off = off; h.data_len = (size_t)bytes;
        off += ((size_t)bytes + 2879) / 2880 * 2880;
        out.push_back(h);
    }
    return !out.empty();
}
static int tform_width(const std::string &tf, char &type) {
    size_t i = 0;
    long long rep = 0;
    while (i < tf.size() && isdigit((unsigned char)tf[i])) rep = rep * 10 + (tf[i++] - '0');
    if (i == 0) rep = 1;
    type = i < tf.size() ? tf[i] : '?';
    int w = 0;
    switch (type) { case 'D': case 'K': case 'M': w = 8; break; case 'E': case 'J': case 'C': w = 4; break;
                    case 'I': w = 2; break; case 'B': case 'L': case 'A': w = 1; break; default: return -1; }
    if (type == 'M') w = 16;
    return (int)(rep * w);
}
// first `want` rows of a float64 ('D') column found by name in any BINTABLE
static bool read_column(const std::vector<unsigned char> &f, const std::vector<Hdu> &hdus, const char *name, int want, double *out) {
    for (auto &h : hdus) {
        if (h.get("XTENSION") != "BINTABLE") continue;
        const long long nf = h.geti("TFIELDS"), rowb = h.geti("NAXIS1"), rows = h.geti("NAXIS2");
        long long coff = 0;
        for (long long i = 1; i <= nf; i++) {
            char type;
            const int w = tform_width(h.get("TFORM" + std::to_string(i)), type);
            if (w < 0) break;
            if (h.get("TTYPE" + std::to_string(i)) == name) {
                if (type != 'D' || rows < want || h.data_off + (size_t)(rows * rowb) > f.size()) return false;
                for (int r = 0; r < want; r++) out[r] = get_be64(&f[h.data_off + (size_t)(r * rowb + coff)]);
                return true;
            }
            coff += w;
        }
    }
    return false;
}


// ---- the full layout ------------------------------------------------------------------------------------------
struct FileOut {
    FILE *fp = nullptr;
    std::string buf;
    bool ok = true;
    size_t total = 0;
    void flush() {
        if (!buf.empty() && ok) ok = fwrite(buf.data(), 1, buf.size(), fp) == buf.size();
        total += buf.size();
        buf.clear();
    }
    void put(const std::string &x) { buf += x; if (buf.size() >= (8u << 20)) flush(); }
    void put_f64(double x) { put_be64(buf, x); if (buf.size() >= (8u << 20)) flush(); }
    void pad() { flush(); std::string z((2880 - total % 2880) % 2880, '\0'); put(z); flush(); }
};
static std::string image_header(const char *extname, bool primary, long long n1, long long n2, long long n3) {
    std::string h;
    if (primary) put_card(h, "SIMPLE", "T");
    else put_card(h, "XTENSION", "IMAGE", true);
    put_card(h, "BITPIX", "-64");
    put_card(h, "NAXIS", "3");
    put_card(h, "NAXIS1", std::to_string(n1));
    put_card(h, "NAXIS2", std::to_string(n2));
    put_card(h, "NAXIS3", std::to_string(n3));
    if (primary) put_card(h, "EXTEND", "T");
    else { put_card(h, "PCOUNT", "0"); put_card(h, "GCOUNT", "1"); }
    put_card(h, "EXTNAME", extname, true);
    end_header(h);
    return h;
}
// columns of differing lengths: rows = the longest, shorter ones zero-filled (astropy FITS_rec.from_columns)
static std::string bintable_ragged(const char *extname, const std::vector<std::pair<std::string, std::pair<const double *, long long>>> &cols) {
    long long rows = 0;
    for (auto &c : cols) rows = std::max(rows, c.second.second);
    std::vector<std::vector<double>> padded(cols.size());
    std::vector<std::pair<std::string, const double *>> flat;
    for (size_t i = 0; i < cols.size(); i++) {
        padded[i].assign((size_t)std::max<long long>(rows, 1), 0.0);
        for (long long r = 0; r < cols[i].second.second; r++) padded[i][(size_t)r] = cols[i].second.first[r];
        flat.push_back({cols[i].first, padded[i].data()});
    }
    return bintable(extname, flat, rows);
}

}  // namespace shopt

using namespace shopt;

extern "C" {

int shopt_summary_write(const char *path, const double *poly, int r, int c, int degree, const double *sC,
                        const double *g1C, const double *g2C) {
    if (!path || !poly || !sC || !g1C || !g2C || r < 1 || c < 1 || degree < 0 || degree > 8) return SHOPT_E_BADARG;
    const int K = (degree + 1) * (degree + 2) / 2;
    const size_t rc = (size_t)r * c;
    std::string h, d;
    put_card(h, "SIMPLE", "T");
    put_card(h, "BITPIX", "-64");
    put_card(h, "NAXIS", "3");
    put_card(h, "NAXIS1", std::to_string(K));
    put_card(h, "NAXIS2", std::to_string(c));
    put_card(h, "NAXIS3", std::to_string(r));
    put_card(h, "EXTEND", "T");
    put_card(h, "EXTNAME", "POLYNOMIAL_MATRIX", true);
    end_header(h);
    d.reserve(rc * K * 8 + 2880);
    for (int a = 0; a < r; a++)                                 // numpy (r, c, K) in C order: K fastest
        for (int b = 0; b < c; b++)
            for (int k = 0; k < K; k++) put_be64(d, poly[(size_t)a + (size_t)b * r + (size_t)k * rc]);
    pad_data(d);
    const double deg = (double)degree;
    const std::string t1 = bintable("SUMMARY_STATISTICS", {{"POLYNOMIAL_DEGREE", &deg}}, 1);
    const std::string t2 = bintable("ANALYTIC_PROFILE", {{"s_MATRIX", sC}, {"g1_MATRIX", g1C}, {"g2_MATRIX", g2C}}, 10);
    FILE *fp = fopen(path, "wb");
    if (!fp) { set_last_error_text("cannot open summary file for writing"); return SHOPT_E_BADARG; }
    const bool ok = fwrite(h.data(), 1, h.size(), fp) == h.size() && fwrite(d.data(), 1, d.size(), fp) == d.size() &&
                    fwrite(t1.data(), 1, t1.size(), fp) == t1.size() && fwrite(t2.data(), 1, t2.size(), fp) == t2.size();
    if (fclose(fp) != 0 || !ok) { set_last_error_text("short write on summary file"); return SHOPT_E_CUDA; }
    return SHOPT_OK;
}

int shopt_summary_write_full(const char *path, const shopt_summary_full *f) {
    if (!path || !f || !f->poly || !f->sC || !f->g1C || !f->g2C || f->r < 1 || f->c < 1 || f->degree < 0 || f->degree > 8)
        return SHOPT_E_BADARG;
    const long long N = f->n_stars, NV = f->n_valid, NT = f->n_train;
    if (N < 0 || NV < 0 || NT < 0 || f->n_flag_snr < 0 || f->n_flag_blacklist < 0 || f->n_flag_failed < 0) return SHOPT_E_BADARG;
    if (N > 0 && (!f->u || !f->v || !f->s_model || !f->g1_model || !f->g2_model || !f->s_data || !f->g1_data || !f->g2_data ||
                  !f->stars || !f->pgf)) return SHOPT_E_BADARG;
    if (NT > 0 && (!f->train_u || !f->train_v)) return SHOPT_E_BADARG;
    if (NV > 0 && (!f->valid_u || !f->valid_v || !f->valid_stars)) return SHOPT_E_BADARG;
    if ((f->n_flag_snr && !f->flag_snr) || (f->n_flag_blacklist && !f->flag_blacklist) || (f->n_flag_failed && !f->flag_failed))
        return SHOPT_E_BADARG;
    const int r = f->r, c = f->c, K = (f->degree + 1) * (f->degree + 2) / 2;
    const size_t rc = (size_t)r * c;
    FileOut out;
    out.fp = fopen(path, "wb");
    if (!out.fp) { set_last_error_text("cannot open summary file for writing"); return SHOPT_E_BADARG; }
    // HDU 0: numpy (r, c, K), K fastest
    out.put(image_header("POLYNOMIAL_MATRIX", true, K, c, r));
    for (int a = 0; a < r; a++)
        for (int b = 0; b < c; b++)
            for (int k = 0; k < K; k++) out.put_f64(f->poly[(size_t)a + (size_t)b * r + (size_t)k * rc]);
    out.pad();
    // HDU 1
    const double deg = (double)f->degree;
    out.put(bintable_ragged("SUMMARY_STATISTICS", {
        {"U_COORDINATES", {f->u, N}}, {"V_COORDINATES", {f->v, N}}, {"TRAINING_U_COORDS", {f->train_u, NT}},
        {"TRAINING_V_COORDS", {f->train_v, NT}}, {"VALIDATION_U_COORDS", {f->valid_u, NV}},
        {"VALIDATION_V_COORDS", {f->valid_v, NV}}, {"S_MODEL", {f->s_model, N}}, {"g1_MODEL", {f->g1_model, N}},
        {"g2_MODEL", {f->g2_model, N}}, {"s_DATA", {f->s_data, N}}, {"g1_DATA", {f->g1_data, N}}, {"g2_DATA", {f->g2_data, N}},
        {"POLYNOMIAL_DEGREE", {&deg, 1}}}));
    out.pad();
    // HDU 2: VIGNETS numpy (r, c, N): star index fastest
    out.put(image_header("VIGNETS", false, N, c, r));
    for (int a = 0; a < r; a++)
        for (int b = 0; b < c; b++)
            for (long long i = 0; i < N; i++) out.put_f64(f->stars[(size_t)a + (size_t)b * r + (size_t)i * rc]);
    out.pad();
    // HDU 3 / 4: numpy (N, r, c): column index fastest
    for (int which = 0; which < 2; which++) {
        const long long n = which ? NV : N;
        const double *cube = which ? f->valid_stars : f->pgf;
        out.put(image_header(which ? "VALIDATION_STARS" : "PIXEL_GRID_FITS", false, c, r, n));
        for (long long i = 0; i < n; i++)
            for (int a = 0; a < r; a++)
                for (int b = 0; b < c; b++) out.put_f64(cube[(size_t)a + (size_t)b * r + (size_t)i * rc]);
        out.pad();
    }
    // HDU 5, 6
    out.put(bintable_ragged("FLAGS", {{"FLAG_SNR", {f->flag_snr, f->n_flag_snr}},
                                      {"FLAG_S_BLACKLIST_MODEL", {f->flag_blacklist, f->n_flag_blacklist}},
                                      {"FLAG_PIXEL_GRID_OR_S_BLACKLIST_LEARNED", {f->flag_failed, f->n_flag_failed}}}));
    out.pad();
    out.put(bintable("ANALYTIC_PROFILE", {{"s_MATRIX", f->sC}, {"g1_MATRIX", f->g1C}, {"g2_MATRIX", f->g2C}}, 10));
    out.pad();
    const bool ok = out.ok;
    if (fclose(out.fp) != 0 || !ok) { set_last_error_text("short write on summary file"); return SHOPT_E_CUDA; }
    return SHOPT_OK;
}

int shopt_summary_read(const char *path, double *poly, int64_t poly_capacity, int *r, int *c, int *degree, double *sC,
                       double *g1C, double *g2C) {
    if (!path || !r || !c || !degree) return SHOPT_E_BADARG;
    FILE *fp = fopen(path, "rb");
    if (!fp) { set_last_error_text("cannot open summary file"); return SHOPT_E_BADARG; }
    std::vector<unsigned char> f;
    unsigned char buf[1 << 16];
    size_t got;
    while ((got = fread(buf, 1, sizeof buf, fp)) > 0) f.insert(f.end(), buf, buf + got);
    fclose(fp);
    std::vector<Hdu> hdus;
    if (!parse_hdus(f, hdus)) { set_last_error_text("not a FITS file"); return SHOPT_E_BADARG; }
    const Hdu &p0 = hdus[0];
    if (p0.geti("NAXIS") != 3 || p0.geti("BITPIX") != -64) { set_last_error_text("primary HDU is not a float64 cube"); return SHOPT_E_BADARG; }
    const long long K = p0.geti("NAXIS1"), cc = p0.geti("NAXIS2"), rr = p0.geti("NAXIS3");
    double deg = -1.0;
    if (!read_column(f, hdus, "POLYNOMIAL_DEGREE", 1, &deg)) { set_last_error_text("no POLYNOMIAL_DEGREE column"); return SHOPT_E_BADARG; }
    *r = (int)rr; *c = (int)cc; *degree = (int)deg;
    if (((long long)*degree + 1) * (*degree + 2) / 2 != K || p0.data_off + (size_t)(K * cc * rr * 8) > f.size()) {
        set_last_error_text("POLYNOMIAL_MATRIX does not match POLYNOMIAL_DEGREE"); return SHOPT_E_BADARG;
    }
    if (poly) {
        if (poly_capacity < K * cc * rr) return SHOPT_E_BADARG;
        const size_t rc = (size_t)rr * cc;
        const unsigned char *src = &f[p0.data_off];
        for (long long a = 0; a < rr; a++)
            for (long long b = 0; b < cc; b++)
                for (long long k = 0; k < K; k++, src += 8) poly[(size_t)a + (size_t)b * rr + (size_t)k * rc] = get_be64(src);
    }
    if (sC && !read_column(f, hdus, "s_MATRIX", 10, sC)) return SHOPT_E_BADARG;
    if (g1C && !read_column(f, hdus, "g1_MATRIX", 10, g1C)) return SHOPT_E_BADARG;
    if (g2C && !read_column(f, hdus, "g2_MATRIX", 10, g2C)) return SHOPT_E_BADARG;
    return SHOPT_OK;
}

}  // extern "C"
