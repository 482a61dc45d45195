// Host-side builder of the B200 tile stream (tile_format.hpp).
#include "tile_format.hpp"

#include <algorithm>
#include <stdexcept>

namespace gs {

namespace {

struct pair16 {
    uint16_t off, val;
};

// One run: place the pairs so that every warp-wide store (pair slot j of the 32 lanes of one warp, 8 slots per lane,
// 8 warps per pass) touches each shared-memory bank at most once where the bank histogram of the tile allows it.
// Greedy: every store takes one pair from each of the fullest banks; only when fewer banks than lanes are left does a
// bank give two.  The sum over the stores of the worst bank multiplicity ends within a few percent of
// max(stores, fullest bank), the lower bound.
void order_run(std::vector<pair16> &run, uint32_t pass_warps)
{
    const uint32_t pass_pairs = pass_warps * 256u;
    const uint32_t n = (uint32_t)run.size(); // multiple of 8
    std::vector<pair16> bank[32];
    for (const pair16 &p : run) bank[(p.off >> 2) & 31].push_back(p);
    std::vector<pair16> out(n);
    int order[32];
    for (uint32_t pass = 0; pass * pass_pairs < n; pass++) {
        for (uint32_t w = 0; w < pass_warps; w++) {
            const uint32_t base = pass * pass_pairs + w * 256u;
            if (base >= n) break;
            for (uint32_t j = 0; j < 8; j++) {
                uint32_t lanes = std::min<uint32_t>(32u, (n - base) / 8u); // lanes whose 8-pair group exists
                for (int b = 0; b < 32; b++) order[b] = b;
                std::sort(order, order + 32, [&](int x, int y) { return bank[x].size() != bank[y].size() ? bank[x].size() > bank[y].size() : x < y; });
                uint32_t l = 0;
                while (l < lanes) {
                    bool any = false;
                    for (int k = 0; k < 32 && l < lanes; k++) {
                        std::vector<pair16> &bk = bank[order[k]];
                        if (bk.empty()) break; // sorted: the rest is empty too
                        out[base + l * 8u + j] = bk.back();
                        bk.pop_back();
                        l++, any = true;
                    }
                    if (!any) throw std::logic_error("tile stream: run ordering ran out of pairs");
                }
            }
        }
    }
    run.swap(out);
}

} // namespace

void build_tile_stream(const uint32_t *rowptr, const uint32_t *col, const uint16_t *val, uint32_t n_rows, uint32_t n_cols, uint32_t pass_warps, tile_stream &ts)
{
    ts = tile_stream();
    ts.n_rows = n_rows, ts.n_cols = n_cols;
    ts.n_panels = (n_rows + TILE_ROWS - 1) / TILE_ROWS;
    ts.k_tiles = (n_cols + TILE_COLS - 1) / TILE_COLS;
    const uint64_t T = (uint64_t)ts.n_panels * ts.k_tiles;
    if (T + 1 > 0xffffffffull) throw std::invalid_argument("tile stream: too many tiles");
    std::vector<uint32_t> cnt(T, 0);
    for (uint32_t r = 0; r < n_rows; r++) {
        const uint64_t p = r / TILE_ROWS;
        for (uint32_t i = rowptr[r]; i < rowptr[r + 1]; i++) {
            if ((val[i] & 0x7fffu) == 0) continue; // +-0: the chain's padding
            if (col[i] >= n_cols) throw std::invalid_argument("tile stream: column index out of range");
            cnt[p * ts.k_tiles + col[i] / TILE_COLS]++;
        }
    }
    ts.tile_ptr.assign(T + 1, 0);
    uint64_t total = 0;
    for (uint64_t t = 0; t < T; t++) {
        ts.tile_ptr[t] = (uint32_t)total;
        ts.nnz += cnt[t];
        total += (cnt[t] + 7u) / 8u * 8u;
        if (total > 0xfffffff0ull) throw std::invalid_argument("tile stream: more than 2^32 pairs");
    }
    ts.tile_ptr[T] = (uint32_t)total;
    ts.off.assign(total, 0), ts.val.assign(total, 0);
    std::vector<uint32_t> fill(T, 0);
    for (uint32_t r = 0; r < n_rows; r++) {
        const uint64_t p = r / TILE_ROWS;
        const uint32_t rr = r % TILE_ROWS;
        for (uint32_t i = rowptr[r]; i < rowptr[r + 1]; i++) {
            if ((val[i] & 0x7fffu) == 0) continue;
            const uint64_t t = p * ts.k_tiles + col[i] / TILE_COLS;
            const uint32_t e = ts.tile_ptr[t] + fill[t]++;
            ts.off[e] = (uint16_t)tile_offset(rr, col[i] % TILE_COLS), ts.val[e] = val[i];
        }
    }
    std::vector<pair16> run;
    for (uint64_t t = 0; t < T; t++) {
        const uint32_t b = ts.tile_ptr[t], e = ts.tile_ptr[t + 1], n = cnt[t];
        if (n == 0) continue;
        if (!pass_warps) {
            for (uint32_t i = b + n; i < e; i++) ts.off[i] = ts.off[b + n - 1], ts.val[i] = ts.val[b + n - 1];
            continue;
        }
        run.resize(e - b);
        for (uint32_t i = 0; i < n; i++) run[i] = {ts.off[b + i], ts.val[b + i]};
        // padding pairs repeat a pair of the emptiest bank that has one: they cost no extra conflict there
        uint32_t hist[32] = {0};
        for (uint32_t i = 0; i < n; i++) hist[(run[i].off >> 2) & 31]++;
        for (uint32_t i = n; i < e - b; i++) {
            int best = -1;
            for (int k = 0; k < 32; k++)
                if (hist[k] && (best < 0 || hist[k] < hist[best])) best = k;
            for (uint32_t q = 0; q < n; q++)
                if ((int)((run[q].off >> 2) & 31) == best) {
                    run[i] = run[q];
                    break;
                }
            hist[best]++;
        }
        order_run(run, pass_warps);
        for (uint32_t i = 0; i < e - b; i++) ts.off[b + i] = run[i].off, ts.val[b + i] = run[i].val;
    }
}

void expand_tile_stream(const tile_stream &ts, std::vector<uint16_t> &dense)
{
    dense.assign((size_t)ts.n_rows * ts.n_cols, 0);
    for (uint64_t t = 0; t + 1 < ts.tile_ptr.size(); t++) {
        const uint32_t p = (uint32_t)(t / ts.k_tiles), kt = (uint32_t)(t % ts.k_tiles);
        for (uint32_t i = ts.tile_ptr[t]; i < ts.tile_ptr[t + 1]; i++) {
            const uint32_t o = ts.off[i], r = o / 128u, chunk = ((o >> 4) & 7u) ^ (r & 7u), c = chunk * 8u + ((o & 15u) >> 1);
            const uint64_t row = (uint64_t)p * TILE_ROWS + r, colx = (uint64_t)kt * TILE_COLS + c;
            if (row < ts.n_rows && colx < ts.n_cols) dense[row * ts.n_cols + colx] = ts.val[i];
        }
    }
}

} // namespace gs
