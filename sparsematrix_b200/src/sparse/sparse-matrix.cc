// sparse-matrix.cc -- host side of sblas::SparseMatrix for the B200 build.
//
// CopyForm / CopyTo stay on the host and produce / consume exactly the byte streams of
// the reference (reference src/sparse/sparse-matrix.cc:21-137; format specification in
// SURVEY.md appendix A).  AddMatMat (reference :139-194) forwards to the CUDA library.
#include "sparse-matrix.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <type_traits>

#include "device_registry.h"
#include "sblas_b200.h"

namespace sblas {

#define SBLAS_TMPL                                                                                          \
    template <typename PosIndex_t, typename ValIndex_t, typename Value_t, const int32 block_row_shift,       \
              const int32 block_col_shift>
#define SBLAS_CLS SparseMatrix<PosIndex_t, ValIndex_t, Value_t, block_row_shift, block_col_shift>

SBLAS_TMPL void SBLAS_CLS::Destroy() {
    device::Evict(this);
    pos_index_.clear();
    val_index_.clear();
    val_table_.clear();
    block_bounds_.clear();
    block_index_bounds_.clear();
    rows_ = cols_ = 0;
}

SBLAS_TMPL SBLAS_CLS &SBLAS_CLS::operator=(const SBLAS_CLS &oth) {
    if (this == &oth) return *this;
    device::Evict(this);                            // the registered device matrix holds the OLD contents
    pos_index_ = oth.pos_index_;
    val_index_ = oth.val_index_;
    val_table_ = oth.val_table_;
    block_bounds_ = oth.block_bounds_;
    block_index_bounds_ = oth.block_index_bounds_;
    rows_ = oth.rows_;
    cols_ = oth.cols_;
    return *this;
}

// CopyForm on the device (src/utils/copyform.cu) for inputs large enough to pay for the trip:
// fills the members with exactly the bytes the scan below produces and registers the device
// matrix built from the device-resident streams, so the first AddMatMat uploads nothing.
// SBLAS_B200_COPYFORM=host|device forces one side (default: device from 2^18 elements).
template <typename Owner>
static bool DeviceCopyForm(Owner *owner, const uint8_t *codes, int32 rows, int32 cols, int32 stride, int32 T, bool trans,
                           const std::vector<float> &val_table, std::vector<uint8_t> *pos, std::vector<uint8_t> *val,
                           std::vector<std::pair<int32, int32> > *bounds, std::vector<std::pair<int32, int32> > *index_bounds) {
    const char *env = getenv("SBLAS_B200_COPYFORM");
    if (env && env[0] == 'h') return false;
    if (!(env && env[0] == 'd') && (int64_t)rows * cols < (1 << 18)) return false;
    const int dev = device::CurrentDevice();
    if (dev < 0) return false;
    sblas_b200_copyform_job *job = nullptr;
    int64_t entries = 0;
    int32 nblocks = 0;
    int rc = sblas_b200_copyform_begin(codes, rows, cols, stride, T, trans ? 1 : 0, dev, &job, &entries, &nblocks);
    if (rc != SBLAS_B200_OK) {
        printf("CopyForm: device path unavailable (%d: %s), scanning on the host\n", rc, sblas_b200_last_error());
        return false;
    }
    pos->resize((size_t)entries);
    val->resize((size_t)entries);
    bounds->resize((size_t)nblocks);
    index_bounds->resize((size_t)nblocks);
    sblas_b200_matrix *mat = nullptr;
    rc = sblas_b200_copyform_finish(job, pos->data(), val->data(), nblocks ? &(*bounds)[0].first : nullptr,
                                    nblocks ? &(*index_bounds)[0].first : nullptr, val_table.data(), &mat);
    if (rc != SBLAS_B200_OK) {
        printf("CopyForm: device path failed (%d: %s), scanning on the host\n", rc, sblas_b200_last_error());
        pos->clear(); val->clear(); bounds->clear(); index_bounds->clear();
        return false;
    }
    device::Adopt(owner, dev, mat);
    return true;
}

template <typename Owner, typename P, typename V, typename F>
static bool DeviceCopyForm(Owner *, const V *, int32, int32, int32, int32, bool, const std::vector<F> &, std::vector<P> *,
                           std::vector<V> *, std::vector<std::pair<int32, int32> > *, std::vector<std::pair<int32, int32> > *) {
    return false;                                   // other instantiations: host scan only
}

// One pass over S in panel order.  S(r, j) is read through `at`, which hides the
// orientation of the dense input; everything else is orientation independent.
SBLAS_TMPL void SBLAS_CLS::CopyForm(const ValIndex_t *density_matrix, int32 rows, int32 cols, int32 stride,
                                    const Value_t *vals, int32 val_table_size, SBLAS_TRANSPOSE trans) {
    static_assert(block_row_shift == 0, "the B200 build implements the reference's only instantiation: full-height panels");
    Destroy();
    const int32 max_delta = (int32)((1u << (8 * sizeof(PosIndex_t))) - 1u);      // 255 for uint8
    const int32 max_code = (int32)((1u << (8 * sizeof(ValIndex_t) < 31 ? 8 * sizeof(ValIndex_t) : 31)) - 1u);   // the filler's code is val_table_size itself
    if (val_table_size < 0 || val_table_size > max_delta || val_table_size > max_code) {
        printf("CopyForm: val_table_size %d outside [0, %d]\n", val_table_size, max_delta < max_code ? max_delta : max_code);
        abort();                                                                   // the reference asserts
    }
    if (val_table_size == 0) return;
    const int32 panel_w = 1 << block_col_shift;
    const bool t = (trans == SblasTrans);
    const int32 s_rows = t ? cols : rows, s_cols = t ? rows : cols;

    val_table_.assign(vals, vals + val_table_size);
    val_table_.push_back(Value_t(0));                                              // trailing zero entry

    if (DeviceCopyForm(this, density_matrix, rows, cols, stride, val_table_size, t, val_table_, &pos_index_, &val_index_,
                       &block_bounds_, &block_index_bounds_)) {
        rows_ = s_rows;
        cols_ = s_cols;
        return;
    }

    for (int32 j0 = 0; j0 < s_cols; j0 += panel_w) {
        const int32 width = (s_cols - j0 < panel_w) ? s_cols - j0 : panel_w;
        const size_t first = pos_index_.size();
        int32 last_lin = 0;
        for (int32 r = 0; r < s_rows; ++r) {
            const ValIndex_t *src = t ? density_matrix + (size_t)j0 * stride + r : density_matrix + (size_t)r * stride + j0;
            const size_t step = t ? (size_t)stride : 1;
            for (int32 jj = 0; jj < width; ++jj, src += step) {
                const ValIndex_t idx = *src;
                if (!(idx >= 0 && idx < val_table_size)) continue;
                const int32 lin = r * panel_w + jj;                                // stride is panel_w even in a narrow panel
                int32 gap = lin - last_lin;
                for (; gap > max_delta; gap -= max_delta) {                        // bridge with filler entries
                    pos_index_.push_back((PosIndex_t)max_delta);
                    val_index_.push_back((ValIndex_t)val_table_size);
                }
                pos_index_.push_back((PosIndex_t)gap);
                val_index_.push_back(idx);
                last_lin = lin;
            }
        }
        if (pos_index_.size() != first) {                                          // empty panels are not recorded
            block_bounds_.push_back(std::make_pair((int32)0, j0));
            block_index_bounds_.push_back(std::make_pair((int32)first, (int32)pos_index_.size()));
        }
    }
    rows_ = s_rows;
    cols_ = s_cols;
}

SBLAS_TMPL void SBLAS_CLS::CopyTo(Value_t *density_matrix, int32 stride, SBLAS_TRANSPOSE trans) {
    const bool t = (trans == SblasTrans);
    const int32 col_mask = (1 << block_col_shift) - 1;
    const int32 filler = (int32)val_table_.size() - 1;
    memset(density_matrix, 0, sizeof(Value_t) * (size_t)(t ? cols_ : rows_) * stride);
    for (size_t b = 0; b < block_bounds_.size(); ++b) {
        const int32 r0 = block_bounds_[b].first, j0 = block_bounds_[b].second;
        int32 lin = 0;
        for (int32 e = block_index_bounds_[b].first; e < block_index_bounds_[b].second; ++e) {
            lin += pos_index_[e];
            if ((int32)val_index_[e] == filler) continue;
            const int32 r = r0 + (lin >> block_col_shift), j = j0 + (lin & col_mask);
            density_matrix[t ? (size_t)j * stride + r : (size_t)r * stride + j] = val_table_[val_index_[e]];
        }
    }
}

SBLAS_TMPL sblas_b200_matrix *SBLAS_CLS::DeviceMatrix() {
    static_assert((std::is_same<PosIndex_t, uint8_t>::value || std::is_same<PosIndex_t, uint16_t>::value) &&
                      std::is_same<ValIndex_t, uint8_t>::value && std::is_same<Value_t, float>::value,
                  "device path exists for <uint8 | uint16, uint8, float>");
    device::StreamView v;
    v.pos_index = pos_index_.data();
    v.pos_bytes = (int32)sizeof(PosIndex_t);
    v.val_index = val_index_.data();
    v.entries = (int64_t)pos_index_.size();
    v.val_table = val_table_.data();
    v.valid_table_size = val_table_.empty() ? 0 : (int32)val_table_.size() - 1;
    v.block_bounds = block_bounds_.empty() ? nullptr : &block_bounds_[0].first;          // pair<int32,int32> is two int32
    v.block_index_bounds = block_index_bounds_.empty() ? nullptr : &block_index_bounds_[0].first;
    v.nblocks = (int32)block_bounds_.size();
    v.rows = rows_;
    v.cols = cols_;
    return device::Acquire(this, v);
}

SBLAS_TMPL void SBLAS_CLS::AddMatMat(Value_t *a, int32 m, int32 lda, Value_t *c, int32 ldc, Value_t alpha, Value_t beta) {
    if (pos_index_.size() != val_index_.size()) {
        printf("AddMatMat: corrupt object (stream lengths differ)\n");
        abort();                                                                   // the reference asserts
    }
    if (cols_ == 0 || m <= 0) return;                                              // empty object: nothing to scale either
    sblas_b200_matrix *dev = DeviceMatrix();
    if (!dev) {
        printf("AddMatMat device matrix unavailable!\n");
        return;
    }
    const int rc = sblas_b200_addmatmat_host(dev, a, m, lda, c, ldc, alpha, beta);
    if (rc != SBLAS_B200_OK) printf("AddMatMat failed (%d): %s\n", rc, sblas_b200_last_error());
}

SBLAS_TMPL bool SBLAS_CLS::operator==(const SBLAS_CLS &oth) {
    return rows_ == oth.rows_ && cols_ == oth.cols_ && pos_index_ == oth.pos_index_ && val_index_ == oth.val_index_ &&
           val_table_ == oth.val_table_ && block_bounds_ == oth.block_bounds_ &&
           block_index_bounds_ == oth.block_index_bounds_;
}

// Known answers of the reference's SelfTest (sparse-matrix.cc:209-313), run through
// the device path: the 3x2 matrix entered in both orientations with the 1x3 product,
// and a 1023x511 round trip in both orientations (deterministic LCG instead of rand()).
SBLAS_TMPL bool SBLAS_CLS::SelfTest() {
    const Value_t table8[8] = {Value_t(1.1), Value_t(2.2), Value_t(3.3), Value_t(4.4),
                               Value_t(5.5), Value_t(6.6), Value_t(7.7), Value_t(8.8)};
    const ValIndex_t z = (ValIndex_t)-1;
    const Value_t want_s[6] = {table8[0], 0, 0, table8[3], table8[7], 0};          // S = [[1.1,0],[0,4.4],[8.8,0]]
    for (int pass = 0; pass < 2; ++pass) {
        const ValIndex_t as_s[6] = {0, z, z, 3, 7, z}, as_st[6] = {0, z, 7, z, 3, z};
        if (pass == 0) CopyForm(as_s, 3, 2, 2, table8, 8, SblasNoTrans);
        else CopyForm(as_st, 2, 3, 3, table8, 8, SblasTrans);
        Value_t out[6] = {1, 1, 1, 1, 1, 1};
        CopyTo(out, 2, SblasNoTrans);
        if (memcmp(out, want_s, sizeof(out)) != 0) return false;
        Value_t a[3] = {Value_t(3.1), 5, 7}, c[2] = {4, 8};
        AddMatMat(a, 1, 3, c, 2, Value_t(1.3), 2);
        if (fabs(c[0] - 92.513) > 1e-3 || fabs(c[1] - 44.6) > 1e-3) return false;
    }
    {
        const int32 m = 1023, n = 511, stride = 512, T = 63;
        std::vector<ValIndex_t> index((size_t)m * stride, z);
        std::vector<Value_t> dense((size_t)m * stride, 0), back((size_t)m * stride, 0), table(T);
        uint64_t lcg = 0x9E3779B97F4A7C15ull;
        auto next = [&lcg]() { lcg = lcg * 6364136223846793005ull + 1442695040888963407ull; return (uint32_t)(lcg >> 33); };
        for (int32 i = 0; i < T; ++i) table[i] = Value_t((int32)(next() % 2001) - 1000) / Value_t(8) + Value_t(0.0625);
        for (int32 i = 0; i < m; ++i)
            for (int32 j = 0; j < n; ++j)
                if (next() % 4 == 0) {
                    index[(size_t)i * stride + j] = (ValIndex_t)(next() % T);
                    dense[(size_t)i * stride + j] = table[index[(size_t)i * stride + j]];
                }
        for (int pass = 0; pass < 2; ++pass) {
            const SBLAS_TRANSPOSE tr = pass ? SblasTrans : SblasNoTrans;
            SBLAS_CLS mat;
            mat.CopyForm(index.data(), m, n, stride, table.data(), T, tr);
            mat.CopyTo(back.data(), stride, tr);
            for (int32 i = 0; i < m; ++i)
                if (memcmp(&dense[(size_t)i * stride], &back[(size_t)i * stride], sizeof(Value_t) * n) != 0) return false;
        }
    }
    return true;
}

template class SparseMatrix<uint8, uint8, float>;
// Two-byte position deltas: the reference has the code path (sparse-matrix.cc:24) but never
// instantiates it (:315).  No fillers below gaps of 65 535: the hypersparse end of the density sweep.
template class SparseMatrix<uint16, uint8, float>;

}  // namespace sblas
