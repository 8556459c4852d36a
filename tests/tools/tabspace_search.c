/* tests/tools/tabspace_search.c — TEST TOOL (SURVEY.md §8 a6'(ii)).
 * Finds the per-length code counts of a COMPLETE prefix code over at most `nsym` symbols (lengths <= 15)
 * whose two-level decoding table, built the way the reference sizes it (root table of 2^root entries plus
 * one sub-table per root prefix that carries longer codes, sized by the longest code under the prefix:
 * /root/reference/src/inftrees.c:222-228,275-297), needs exactly `target` entries.  Exhaustive search over
 * the count vectors with memoisation of the failed states (the approach of zlib's examples/enough.c, written
 * from the description above).  usage: tabspace_search nsym root target  ->  prints c[1..15]
 *
 * Forward accounting: while codes of length l > root are placed, an "open" sub-table has size 2^(l-root) and
 * `rem` unfilled slots; moving on to length l+1 with rem > 0 doubles it (mem += 2^(l-root), rem *= 2). */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

static int nsym, root, target, cnt[16];
#define HBITS 26
static uint64_t *seen;
static int seen_put(uint64_t k) {          /* 1 = already there */
    uint64_t h = (k * 0x9E3779B97F4A7C15ull) >> (64 - HBITS);
    for (;;) {
        if (seen[h] == k) return 1;
        if (seen[h] == 0) { seen[h] = k; return 0; }
        h = (h + 1) & ((1u << HBITS) - 1);
    }
}
/* l = length whose count is chosen next; left = unused patterns at length l; syms = symbols left */
static int dfs(int l, int left, int syms, int mem, int rem) {
    if (left == 0) return mem == target && rem == 0;
    if (l > 15 || syms == 0 || mem > target) return 0;
    uint64_t key = 1 + (uint64_t)l + 16ull * left + 16ull * 65536 * syms + 16ull * 65536 * 512 * mem + 16ull * 65536 * 512 * 1024 * (uint64_t)rem;
    if (seen_put(key)) return 0;
    int cmax = left < syms ? left : syms;
    for (int c = cmax; c >= 0; c--) {
        if (l == 15 && c != left) continue;
        int m = mem, r = rem;
        if (l > root) {
            int sz = 1 << (l - root), cc = c;
            int use = cc < r ? cc : r;
            r -= use; cc -= use;
            if (cc > 0) {                      /* r == 0 here: new sub-tables */
                int tabs = (cc + sz - 1) / sz;
                m += tabs * sz;
                r = tabs * sz - cc;
            }
            if (r > 0 && left - c > 0) {       /* the open sub-table continues at the next length */
                m += sz;
                r *= 2;
            }
        }
        cnt[l] = c;
        if (dfs(l + 1, (left - c) * 2, syms - c, m, r)) return 1;
    }
    cnt[l] = 0;
    return 0;
}
int main(int argc, char **argv) {
    if (argc < 4) return 2;
    nsym = atoi(argv[1]); root = atoi(argv[2]); target = atoi(argv[3]);
    seen = calloc((size_t)1 << HBITS, sizeof *seen);
    if (dfs(1, 2, nsym, 1 << root, 0)) {
        printf("[0");
        for (int l = 1; l <= 15; l++) printf(", %d", cnt[l]);
        printf("]\n");
        return 0;
    }
    printf("None\n");
    return 1;
}
