// host_capi.cc -- extern "C" handle API over sblas::SparseMatrix (include/sblas_b200_host.h).
// This translation unit is the class's only friend-less client that needs the compressed
// members, so it reads them through a layout-identical view (the members are private, as in
// the reference, sparse-matrix.h:45-52).
#include <stdio.h>
#include <string.h>

#include <new>
#include <vector>

#include "sblas_b200_host.h"
#include "sparse-matrix.h"

typedef sblas::SparseMatrix<uint8, uint8, float> HostMatrix;

namespace {
// Same member sequence as HostMatrix; checked by size below and by the structure parity tests.
struct MemberView {
    std::vector<uint8_t> pos_index;
    std::vector<uint8_t> val_index;
    std::vector<float> val_table;
    std::vector<std::pair<int32, int32> > block_bounds;
    std::vector<std::pair<int32, int32> > block_index_bounds;
    int32 rows, cols;
};
static_assert(sizeof(MemberView) == sizeof(HostMatrix), "SparseMatrix layout changed");
inline const MemberView &view(const sblas_host_matrix *h) { return *reinterpret_cast<const MemberView *>(h); }
inline HostMatrix *obj(sblas_host_matrix *h) { return reinterpret_cast<HostMatrix *>(h); }
inline const HostMatrix *obj(const sblas_host_matrix *h) { return reinterpret_cast<const HostMatrix *>(h); }
inline MemberView &mutable_view(sblas_host_matrix *h) { return *reinterpret_cast<MemberView *>(h); }

const char kMagic[8] = {'S', 'B', 'L', 'A', 'S', 'M', '1', '\0'};
struct FileHeader { char magic[8]; int32_t rows, cols, table_size, nblocks; int64_t entries; };

uint64_t fnv1a(uint64_t hsh, const void *data, size_t bytes) {
    const unsigned char *p = static_cast<const unsigned char *>(data);
    for (size_t i = 0; i < bytes; ++i) { hsh ^= p[i]; hsh *= 1099511628211ull; }
    return hsh;
}
struct Writer {
    FILE *f; uint64_t hsh; bool ok;
    void put(const void *data, size_t bytes) {
        if (bytes == 0) return;
        hsh = fnv1a(hsh, data, bytes);
        ok = ok && fwrite(data, 1, bytes, f) == bytes;
    }
};
}  // namespace

extern "C" {

sblas_host_matrix *sblas_host_create(void) { return reinterpret_cast<sblas_host_matrix *>(new (std::nothrow) HostMatrix()); }
void sblas_host_free(sblas_host_matrix *h) { delete obj(h); }
void sblas_host_destroy(sblas_host_matrix *h) { if (h) obj(h)->Destroy(); }

int sblas_host_copyform(sblas_host_matrix *h, const uint8_t *density_matrix, int32_t rows, int32_t cols, int32_t stride,
                        const float *vals, int32_t val_table_size, int32_t trans) {
    if (!h || val_table_size < 0 || val_table_size > 255) return SBLAS_B200_EINVAL;   // the class aborts like the reference asserts
    obj(h)->CopyForm(density_matrix, rows, cols, stride, vals, val_table_size, trans ? sblas::SblasTrans : sblas::SblasNoTrans);
    return SBLAS_B200_OK;
}
void sblas_host_copyto(sblas_host_matrix *h, float *density_matrix, int32_t stride, int32_t trans) {
    obj(h)->CopyTo(density_matrix, stride, trans ? sblas::SblasTrans : sblas::SblasNoTrans);
}
int sblas_host_addmatmat(sblas_host_matrix *h, const float *a, int32_t m, int32_t lda, float *c, int32_t ldc, float alpha,
                         float beta) {
    if (!h) return SBLAS_B200_EINVAL;
    if (obj(h)->NumCols() == 0 || m <= 0) return SBLAS_B200_OK;
    sblas_b200_matrix *dev = obj(h)->DeviceMatrix();
    if (!dev) return SBLAS_B200_ECUDA;
    return sblas_b200_addmatmat_host(dev, a, m, lda, c, ldc, alpha, beta);
}
int32_t sblas_host_num_rows(const sblas_host_matrix *h) { return obj(h)->NumRows(); }
int32_t sblas_host_num_cols(const sblas_host_matrix *h) { return obj(h)->NumCols(); }
int32_t sblas_host_equal(sblas_host_matrix *a, sblas_host_matrix *b) { return (*obj(a) == *obj(b)) ? 1 : 0; }
int32_t sblas_host_selftest(void) { HostMatrix t; return t.SelfTest() ? 1 : 0; }

int64_t sblas_host_entries(const sblas_host_matrix *h) { return (int64_t)view(h).pos_index.size(); }
int32_t sblas_host_nblocks(const sblas_host_matrix *h) { return (int32_t)view(h).block_bounds.size(); }
int32_t sblas_host_table_size(const sblas_host_matrix *h) {
    return view(h).val_table.empty() ? 0 : (int32_t)view(h).val_table.size() - 1;
}
void sblas_host_members(const sblas_host_matrix *h, const uint8_t **pos_index, const uint8_t **val_index,
                        const float **val_table, const int32_t **block_bounds, const int32_t **block_index_bounds) {
    const MemberView &v = view(h);
    if (pos_index) *pos_index = v.pos_index.data();
    if (val_index) *val_index = v.val_index.data();
    if (val_table) *val_table = v.val_table.data();
    if (block_bounds) *block_bounds = v.block_bounds.empty() ? nullptr : &v.block_bounds[0].first;
    if (block_index_bounds) *block_index_bounds = v.block_index_bounds.empty() ? nullptr : &v.block_index_bounds[0].first;
}
sblas_b200_matrix *sblas_host_device_matrix(sblas_host_matrix *h) { return h ? obj(h)->DeviceMatrix() : nullptr; }

int sblas_host_save(const sblas_host_matrix *h, const char *path) {
    if (!h || !path) return SBLAS_B200_EINVAL;
    const MemberView &v = view(h);
    if (v.pos_index.size() != v.val_index.size() || v.block_bounds.size() != v.block_index_bounds.size()) return SBLAS_B200_EFORMAT;
    FILE *f = fopen(path, "wb");
    if (!f) return SBLAS_B200_EIO;
    FileHeader hd;
    memcpy(hd.magic, kMagic, 8);
    hd.rows = v.rows; hd.cols = v.cols;
    hd.table_size = v.val_table.empty() ? 0 : (int32_t)v.val_table.size() - 1;
    hd.nblocks = (int32_t)v.block_bounds.size();
    hd.entries = (int64_t)v.pos_index.size();
    Writer w = {f, 14695981039346656037ull, true};
    w.put(&hd, sizeof(hd));
    w.put(v.val_table.data(), v.val_table.size() * sizeof(float));
    w.put(v.block_bounds.data(), v.block_bounds.size() * 8);
    w.put(v.block_index_bounds.data(), v.block_index_bounds.size() * 8);
    w.put(v.pos_index.data(), v.pos_index.size());
    w.put(v.val_index.data(), v.val_index.size());
    const uint64_t sum = w.hsh;
    bool ok = w.ok && fwrite(&sum, 1, 8, f) == 8;
    ok = (fclose(f) == 0) && ok;
    return ok ? SBLAS_B200_OK : SBLAS_B200_EIO;
}

int sblas_host_load(sblas_host_matrix *h, const char *path) {
    if (!h || !path) return SBLAS_B200_EINVAL;
    FILE *f = fopen(path, "rb");
    if (!f) return SBLAS_B200_EIO;
    FileHeader hd;
    MemberView in;                                        // filled and validated first; the object changes only on success
    uint64_t hsh = 14695981039346656037ull, sum = 0;
    int rc = SBLAS_B200_OK;
    auto get = [&](void *data, size_t bytes) {
        if (bytes == 0) return true;
        if (fread(data, 1, bytes, f) != bytes) return false;
        hsh = fnv1a(hsh, data, bytes);
        return true;
    };
    if (!get(&hd, sizeof(hd))) rc = SBLAS_B200_EFORMAT;
    if (rc == SBLAS_B200_OK && (memcmp(hd.magic, kMagic, 8) != 0 || hd.rows < 0 || hd.cols < 0 || hd.table_size < 0 ||
                                hd.table_size > 255 || hd.nblocks < 0 || hd.entries < 0 || hd.entries >= ((int64_t)1 << 31) ||
                                hd.nblocks > (hd.cols + 255) / 256 || (hd.table_size == 0 && (hd.entries || hd.nblocks))))
        rc = SBLAS_B200_EFORMAT;
    if (rc == SBLAS_B200_OK) {
        try {
            in.val_table.resize(hd.table_size ? (size_t)hd.table_size + 1 : 0);
            in.block_bounds.resize((size_t)hd.nblocks);
            in.block_index_bounds.resize((size_t)hd.nblocks);
            in.pos_index.resize((size_t)hd.entries);
            in.val_index.resize((size_t)hd.entries);
        } catch (...) { rc = SBLAS_B200_ENOMEM; }
    }
    if (rc == SBLAS_B200_OK &&
        !(get(in.val_table.data(), in.val_table.size() * sizeof(float)) && get(in.block_bounds.data(), in.block_bounds.size() * 8) &&
          get(in.block_index_bounds.data(), in.block_index_bounds.size() * 8) && get(in.pos_index.data(), in.pos_index.size()) &&
          get(in.val_index.data(), in.val_index.size()) && fread(&sum, 1, 8, f) == 8 && sum == hsh && fgetc(f) == EOF))
        rc = SBLAS_B200_EFORMAT;
    fclose(f);
    if (rc != SBLAS_B200_OK) return rc;
    int32_t expect = 0, last_col = -256;
    for (int32_t b = 0; b < hd.nblocks; ++b) {            // panels ascending, streams contiguous (sparse-matrix.cc:35,57-60)
        const int32_t col_off = in.block_bounds[b].second, s = in.block_index_bounds[b].first, e = in.block_index_bounds[b].second;
        if (in.block_bounds[b].first != 0 || (col_off & 255) || col_off <= last_col || col_off >= hd.cols || s != expect || e <= s)
            return SBLAS_B200_EFORMAT;
        expect = e; last_col = col_off;
    }
    if ((int64_t)expect != hd.entries) return SBLAS_B200_EFORMAT;
    for (size_t i = 0; i < in.val_index.size(); ++i)
        if ((int32_t)in.val_index[i] > hd.table_size) return SBLAS_B200_EFORMAT;
    // every stored entry must decode to a position inside the matrix (CopyTo has no bounds check, like the reference:
    // a crafted file with a matching checksum would otherwise write outside the caller's dense buffer), and a real
    // entry after the first of a panel has a delta >= 1 (two entries on one position would collapse in the dense-code SpMV)
    for (int32_t b = 0; b < hd.nblocks; ++b) {
        const int32_t col_off = in.block_bounds[b].second;
        const int32_t width = hd.cols - col_off < 256 ? hd.cols - col_off : 256;
        int64_t lin = 0;
        bool first = true;
        for (int32_t e = in.block_index_bounds[b].first; e < in.block_index_bounds[b].second; ++e) {
            lin += in.pos_index[e];
            if ((int32_t)in.val_index[e] == hd.table_size) continue;          // filler
            if ((lin >> 8) >= hd.rows || (int32_t)(lin & 255) >= width) return SBLAS_B200_EFORMAT;
            if (!first && in.pos_index[e] == 0) return SBLAS_B200_EFORMAT;
            first = false;
        }
    }
    obj(h)->Destroy();                                    // drops the device copy, like CopyForm (sparse-matrix.cc:22)
    MemberView &out = mutable_view(h);
    out.pos_index.swap(in.pos_index);
    out.val_index.swap(in.val_index);
    out.val_table.swap(in.val_table);
    out.block_bounds.swap(in.block_bounds);
    out.block_index_bounds.swap(in.block_index_bounds);
    out.rows = hd.rows;
    out.cols = hd.cols;
    return SBLAS_B200_OK;
}


/* ---- sblas::SparseMatrix<uint16, uint8, float>: two-byte position deltas ---------------------------------------- */
typedef sblas::SparseMatrix<uint16, uint8, float> HostMatrix16;
namespace {
struct MemberView16 {
    std::vector<uint16_t> pos_index;
    std::vector<uint8_t> val_index;
    std::vector<float> val_table;
    std::vector<std::pair<int32, int32> > block_bounds;
    std::vector<std::pair<int32, int32> > block_index_bounds;
    int32 rows, cols;
};
static_assert(sizeof(MemberView16) == sizeof(HostMatrix16), "SparseMatrix layout changed");
inline HostMatrix16 *obj16(sblas_host_matrix16 *h) { return reinterpret_cast<HostMatrix16 *>(h); }
inline const MemberView16 &view16(const sblas_host_matrix16 *h) { return *reinterpret_cast<const MemberView16 *>(h); }
}  // namespace

sblas_host_matrix16 *sblas_host16_create(void) { return reinterpret_cast<sblas_host_matrix16 *>(new (std::nothrow) HostMatrix16()); }
void sblas_host16_free(sblas_host_matrix16 *h) { delete obj16(h); }
int sblas_host16_copyform(sblas_host_matrix16 *h, const uint8_t *density_matrix, int32_t rows, int32_t cols, int32_t stride,
                          const float *vals, int32_t val_table_size, int32_t trans) {
    if (!h || val_table_size < 0 || val_table_size > 255) return SBLAS_B200_EINVAL;
    obj16(h)->CopyForm(density_matrix, rows, cols, stride, vals, val_table_size, trans ? sblas::SblasTrans : sblas::SblasNoTrans);
    return SBLAS_B200_OK;
}
void sblas_host16_copyto(sblas_host_matrix16 *h, float *density_matrix, int32_t stride, int32_t trans) {
    obj16(h)->CopyTo(density_matrix, stride, trans ? sblas::SblasTrans : sblas::SblasNoTrans);
}
int sblas_host16_addmatmat(sblas_host_matrix16 *h, const float *a, int32_t m, int32_t lda, float *c, int32_t ldc, float alpha,
                           float beta) {
    if (!h) return SBLAS_B200_EINVAL;
    if (obj16(h)->NumCols() == 0 || m <= 0) return SBLAS_B200_OK;
    sblas_b200_matrix *dev = obj16(h)->DeviceMatrix();
    if (!dev) return SBLAS_B200_ECUDA;
    return sblas_b200_addmatmat_host(dev, a, m, lda, c, ldc, alpha, beta);
}
int32_t sblas_host16_num_rows(const sblas_host_matrix16 *h) { return reinterpret_cast<const HostMatrix16 *>(h)->NumRows(); }
int32_t sblas_host16_num_cols(const sblas_host_matrix16 *h) { return reinterpret_cast<const HostMatrix16 *>(h)->NumCols(); }
int64_t sblas_host16_entries(const sblas_host_matrix16 *h) { return (int64_t)view16(h).pos_index.size(); }
int32_t sblas_host16_nblocks(const sblas_host_matrix16 *h) { return (int32_t)view16(h).block_bounds.size(); }
int32_t sblas_host16_table_size(const sblas_host_matrix16 *h) {
    return view16(h).val_table.empty() ? 0 : (int32_t)view16(h).val_table.size() - 1;
}
void sblas_host16_members(const sblas_host_matrix16 *h, const uint16_t **pos_index, const uint8_t **val_index, const float **val_table,
                          const int32_t **block_bounds, const int32_t **block_index_bounds) {
    const MemberView16 &v = view16(h);
    if (pos_index) *pos_index = v.pos_index.data();
    if (val_index) *val_index = v.val_index.data();
    if (val_table) *val_table = v.val_table.data();
    if (block_bounds) *block_bounds = v.block_bounds.empty() ? nullptr : &v.block_bounds[0].first;
    if (block_index_bounds) *block_index_bounds = v.block_index_bounds.empty() ? nullptr : &v.block_index_bounds[0].first;
}
sblas_b200_matrix *sblas_host16_device_matrix(sblas_host_matrix16 *h) { return h ? obj16(h)->DeviceMatrix() : nullptr; }

}  // extern "C"
