/*
 * brisk_oracle.c — CPU restatement of Brisk's k-mer insertion path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this file's library, and only as the checker.
 *
 * Parity status: PINNED.  Every function below restates a reference function
 * (cited file:line, relative to the reference checkout) and is checked in
 * tests/test_oracle_golden.py against fixtures produced by running the
 * unmodified reference in oracle/_ref (see oracle/Makefile, oracle/ref_dump.cpp,
 * tests/golden/make_golden.py).
 *
 * One deliberate deviation: k = 64.  The reference's k_mask is ((kint)1 << 128) - 1
 * (brisk/Kmers.cpp:512), which is undefined and evaluates to 0 with gcc on x86-64.
 * The oracle uses the all-ones mask for k >= 64, the same one-line fix that
 * oracle/Makefile applies to the k64 build of the reference.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef unsigned __int128 kint;

typedef struct {
    int k, m, b;
    int m_reduc;        /* parameters.hpp:28 */
    int s;              /* (m_reduc + 1) / 2, Brisk.hpp:135 */
    kint k_mask;
    uint64_t m_mask;
    double coef[4 * 33];
    double eps;
} oracle_params;

/* ---------------------------------------------------------------- Decycling.cpp */

/* DecyclingSet::DecyclingSet, brisk/Decycling.cpp:7-13 (here k == m) */
static void dec_init(oracle_params *p) {
    int k = p->m;
    double unit = 2 * M_PI / k;
    memset(p->coef, 0, sizeof(p->coef));
    for (size_t i = 4; i < (size_t)4 * k; i += 4) {
        p->coef[i + 1] = sin(unit * (i / 4));
        p->coef[i + 2] = 2 * p->coef[i + 1];
        p->coef[i + 3] = 3 * p->coef[i + 1];
    }
    p->eps = 0.000001; /* Decycling.h:19 */
}

/* DecyclingSet::computeR, brisk/Decycling.cpp:17-24 */
static double dec_computeR(const oracle_params *p, uint64_t seq) {
    double R = 0;
    for (size_t i = 4 * ((size_t)p->m - 1); i > 0; i -= 4) {
        R += p->coef[i + (seq & 3)];
        seq >>= 2;
    }
    return R;
}

/* DecyclingSet::memDouble, brisk/Decycling.cpp:38-52 */
static unsigned dec_memDouble(const oracle_params *p, uint64_t seq) {
    int k = p->m;
    double Rseq = dec_computeR(p, seq);
    if (Rseq > p->eps) {
        uint64_t rot = ((seq & 3) << (2 * (k - 1))) + (seq >> 2);
        if (dec_computeR(p, rot) < p->eps) return 0;
    } else if (Rseq < -p->eps) {
        uint64_t rot = ((seq & 3) << (2 * (k - 1))) + (seq >> 2);
        if (dec_computeR(p, rot) > -p->eps) return 1;
    }
    return 2;
}

/* ------------------------------------------------------------------ hashing.cpp */

/* the integer mix of bfc_hash_64 without the class bits, brisk/hashing.cpp:10-16 */
static uint64_t bfc_mix(uint64_t key, uint64_t mask) {
    key = (~key + (key << 21)) & mask;
    key = key ^ key >> 24;
    key = ((key + (key << 3)) + (key << 8)) & mask;
    key = key ^ key >> 14;
    key = ((key + (key << 2)) + (key << 4)) & mask;
    key = key ^ key >> 28;
    key = (key + (key << 31)) & mask;
    return key;
}

/* bfc_hash_64, brisk/hashing.cpp:8-19 */
static uint64_t bfc_hash_64(const oracle_params *p, uint64_t key, uint64_t mask) {
    uint64_t heavy = dec_memDouble(p, key);
    return (heavy << 62) + bfc_mix(key, mask);
}

/* bfc_hash_64_inv, brisk/hashing.cpp:23-49 */
static uint64_t bfc_hash_64_inv(uint64_t key, uint64_t mask) {
    uint64_t tmp;
    tmp = (key - (key << 31));
    key = (key - (tmp << 31)) & mask;
    tmp = key ^ key >> 28;
    key = key ^ tmp >> 28;
    key = (key * 14933078535860113213ull) & mask;
    tmp = key ^ key >> 14;
    tmp = key ^ tmp >> 14;
    tmp = key ^ tmp >> 14;
    key = key ^ tmp >> 14;
    key = (key * 15244667743933553977ull) & mask;
    tmp = key ^ key >> 24;
    key = key ^ tmp >> 24;
    tmp = ~key;
    tmp = ~(key - (tmp << 21));
    tmp = ~(key - (tmp << 21));
    key = ~(key - (tmp << 21)) & mask;
    return key;
}

/* -------------------------------------------------------------------- Kmers.cpp */

/* rcbc, brisk/Kmers.cpp:320-332: true reverse complement of an n-mer, n <= 32 */
static uint64_t rcbc(uint64_t in, uint64_t n) {
    uint64_t res = __builtin_bswap64(in ^ 0xaaaaaaaaaaaaaaaaull);
    const uint64_t c1 = 0x0f0f0f0f0f0f0f0full;
    const uint64_t c2 = 0x3333333333333333ull;
    res = ((res & c1) << 4) | ((res & (c1 << 4)) >> 4);
    res = ((res & c2) << 2) | ((res & (c2 << 2)) >> 2);
    res >>= 64 - 2 * n;
    return res;
}

/* canonize(uint64_t), brisk/Kmers.cpp:336-338 */
static uint64_t canonize64(uint64_t x, uint64_t n) {
    uint64_t r = rcbc(x, n);
    return x < r ? x : r;
}

/* rcb, brisk/Kmers.cpp:293-316.  The byte shuffle's result is discarded in the
 * reference (line 304), so bytes are NOT reversed: each u64 half only gets its
 * nucleotides swapped inside every byte and complemented, then the whole 128-bit
 * value is shifted right by 128 - 2n.  Restated as written. */
static kint rcb(kint in, uint64_t n) {
    uint64_t h[2] = {(uint64_t)in, (uint64_t)(in >> 64)};
    const uint64_t c1 = 0x0f0f0f0f0f0f0f0full;
    const uint64_t c2 = 0x3333333333333333ull;
    for (int i = 0; i < 2; i++) {
        uint64_t x = h[i];
        x = ((x & c1) << 4) | ((x & (c1 << 4)) >> 4);
        x = ((x & c2) << 2) | ((x & (c2 << 2)) >> 2);
        x ^= 0xaaaaaaaaaaaaaaaaull;
        h[i] = x;
    }
    kint res = ((kint)h[1] << 64) | h[0];
    unsigned count = 128 - 2 * (unsigned)n; /* mm_bitshift_right, Kmers.cpp:282-290 */
    if (count >= 128) return 0;
    if (count == 0) return res;
    return res >> count;
}

/* canonized(kint), brisk/Kmers.cpp:342-353 */
static int canonized128(kint x, uint64_t n) {
    kint r = rcb(x, n);
    kint c = x < r ? x : r;
    return x == c;
}

/* get_minimizer, brisk/Kmers.cpp:367-408 — literal, including the 64-bit
 * truncation of the k-mer (line 371/378-379) and the tie rules
 * (lines 389-404, Q5). */
static uint64_t get_minimizer(const oracle_params *p, kint seq, const uint8_t k,
                              uint8_t *min_position, const uint8_t m, int *reversed,
                              const uint64_t m_mask) {
    uint64_t mini, mmer, hash_mini, new_hash;
    uint64_t fwd_mmer = (uint64_t)(seq & m_mask);
    uint64_t cur_seq = (uint64_t)seq;
    mini = canonize64(fwd_mmer, m);
    hash_mini = bfc_hash_64(p, mini, m_mask);
    *reversed = (mini != fwd_mmer);
    *min_position = 0;
    for (uint64_t i = 1; i <= (unsigned)k - m; i++) {
        cur_seq >>= 2;
        fwd_mmer = cur_seq & m_mask;
        mmer = canonize64(fwd_mmer, m);
        new_hash = bfc_hash_64(p, mmer, m_mask);
        if (new_hash < hash_mini) {
            *min_position = (uint8_t)i;
            mini = mmer;
            *reversed = (mini != fwd_mmer);
            hash_mini = new_hash;
        } else if (hash_mini == new_hash) {
            if ((uint64_t)(k - m) - i < *min_position) {
                *min_position = (uint8_t)(k - m - i);
                mini = mmer;
                *reversed = (mini != fwd_mmer);
                hash_mini = new_hash;
            } else if ((uint64_t)(k - m) - i == *min_position) {
                if (!canonized128(seq, k)) {
                    *min_position = (uint8_t)(k - m - i);
                    mini = mmer;
                    *reversed = 0;
                }
            }
        }
    }
    return mini;
}

/* nuc2int, brisk/Kmers.cpp:442-444: A=0 C=1 T=2 G=3 */
static uint64_t nuc2int(char c) { return (c >> 1) & 3; }

/* updateK / updateRCK, brisk/Kmers.cpp:454-473 */
static void updateK(kint *min, uint8_t nuc, kint mask) {
    *min <<= 2;
    *min += nuc;
    *min &= mask;
}
static void updateRCK(kint *min, char nuc, uint8_t k) {
    *min >>= 2;
    *min += ((kint)nuc << (2 * k - 2));
}

/* init_kmer, brisk/Kmers.cpp:487-496 */
static void init_kmer(const char *seq, kint *kmer_seq, kint *rc_kmer_seq, const uint8_t k,
                      const kint k_mask) {
    *kmer_seq = 0;
    *rc_kmer_seq = 0;
    for (uint8_t i = 0; i < k; i++) {
        uint64_t nuc = nuc2int(seq[i]);
        updateK(kmer_seq, (uint8_t)nuc, k_mask);
        nuc = nuc ^ 2;
        updateRCK(rc_kmer_seq, (char)nuc, k);
    }
}

/* update_kmer, brisk/Kmers.cpp:500-505 */
static void update_kmer(const char nucl, kint *kmer_seq, kint *rc_kmer_seq, const uint8_t k,
                        const kint k_mask) {
    uint64_t nuc = nuc2int(nucl);
    updateK(kmer_seq, (uint8_t)nuc, k_mask);
    nuc = nuc ^ 2;
    updateRCK(rc_kmer_seq, (char)nuc, k);
}

/* kmer_full, brisk/Kmers.hpp:32-58 — only the fields the path uses */
typedef struct {
    kint kmer_s;
    kint minimizer;
    uint8_t minimizer_idx;
    /* instrumentation only (not in the reference): which read k-mer this is,
     * and the enumerator state it was emitted under */
    uint32_t ridx;
    uint8_t rev, pos;
} okmer;

/* kmer_full::kmer_full(value, idx, msize), brisk/Kmers.cpp:17-25 */
static okmer okmer_make(kint value, uint8_t minimizer_idx, uint8_t minimizer_size) {
    okmer r;
    r.minimizer_idx = minimizer_idx;
    r.kmer_s = value;
    r.minimizer = value >> (2 * minimizer_idx);
    r.minimizer &= ((kint)1 << (2 * minimizer_size)) - 1;
    r.ridx = 0;
    r.rev = 0;
    r.pos = 0;
    return r;
}

/* growable vector<kmer_full> stand-in */
typedef struct {
    okmer *v;
    size_t n, cap;
} okvec;
static void okvec_push(okvec *a, okmer x) {
    if (a->n == a->cap) {
        a->cap = a->cap ? 2 * a->cap : 64;
        a->v = (okmer *)realloc(a->v, a->cap * sizeof(okmer));
    }
    a->v[a->n++] = x;
}
static void okvec_reverse(okvec *a) {
    for (size_t i = 0, j = a->n; i + 1 < j; i++) {
        j--;
        okmer t = a->v[i];
        a->v[i] = a->v[j];
        a->v[j] = t;
    }
}

/* SuperKmerEnumerator, brisk/Kmers.hpp:62-89, ctor brisk/Kmers.cpp:509-518 */
typedef struct {
    const oracle_params *p;
    const char *seq;
    uint64_t len;
    uint64_t seq_idx;
    uint8_t k, m;
    kint k_mask, m_mask;
    int saved;
    okmer saved_kmer;
    kint current_kmer, current_rc_kmer;
    kint mini_candidate, rc_mini_candidate;
    int reversed;
    uint8_t mini_pos;
    uint64_t mini, mini_hash;
} oenum;

static void oenum_init(oenum *e, const oracle_params *p, const char *seq, uint64_t len) {
    memset(e, 0, sizeof(*e));
    e->p = p;
    e->seq = seq;
    e->len = len;
    e->seq_idx = 0;
    e->k = (uint8_t)p->k;
    e->m = (uint8_t)p->m;
    e->k_mask = p->k_mask;
    e->m_mask = ((kint)1 << (2 * p->m)) - 1;
    e->saved = 0;
    e->saved_kmer = okmer_make(0, 0, 0);
    e->mini_pos = 1;
    e->mini_hash = 0;
    /* `reversed` and `mini` are left uninitialised by the reference ctor; every
     * path writes them in the seq_idx == 0 block of next() before they are read. */
}

/* SuperKmerEnumerator::next, brisk/Kmers.cpp:522-603 — literal state machine
 * (SURVEY.md Q4).  Appends the next superkmer's k-mers to `kmers` and returns the
 * minimizer value, exactly as the reference does. */
static kint oenum_next(oenum *e, okvec *kmers) {
    const oracle_params *p = e->p;
    const uint8_t k = e->k, m = e->m;
    int to_return = 0;
    uint64_t return_val = 0;
    if (e->seq_idx == 0) {
        init_kmer(e->seq, &e->current_kmer, &e->current_rc_kmer, k - 1, e->k_mask >> 2);
        e->current_rc_kmer <<= 2;
        /* init_kmer(seq.substr(k-m-1, m), ..., m, m_mask>>2), Kmers.cpp:531 */
        init_kmer(e->seq + (k - m - 1), &e->mini_candidate, &e->rc_mini_candidate, m,
                  e->m_mask >> 2);
        e->mini = get_minimizer(p, e->current_kmer, k - 1, &e->mini_pos, m, &e->reversed,
                                (uint64_t)e->m_mask);
        e->mini_hash = bfc_hash_64(p, e->mini, (uint64_t)e->m_mask);
    }
    if (e->saved) {
        e->saved = 0;
        okvec_push(kmers, e->saved_kmer);
    }
    uint64_t line_size = e->len;
    for (; k - 1 + e->seq_idx < line_size; ++e->seq_idx) {
        update_kmer(e->seq[k - 1 + e->seq_idx], &e->current_kmer, &e->current_rc_kmer, k,
                    e->k_mask);
        update_kmer(e->seq[k - 1 + e->seq_idx], &e->mini_candidate, &e->rc_mini_candidate, m,
                    e->m_mask);
        e->mini_pos += 1;
        kint candidate_canon =
            e->mini_candidate < e->rc_mini_candidate ? e->mini_candidate : e->rc_mini_candidate;
        uint64_t current_hash = bfc_hash_64(p, (uint64_t)candidate_canon, (uint64_t)e->m_mask);
        if (e->mini_pos > k - m) {
            if (e->reversed) okvec_reverse(kmers);
            to_return = 1;
            return_val = e->mini;
            e->mini = get_minimizer(p, e->current_kmer, k, &e->mini_pos, m, &e->reversed,
                                    (uint64_t)e->m_mask);
            e->mini_hash = bfc_hash_64(p, e->mini, (uint64_t)e->m_mask);
        } else if (current_hash < e->mini_hash) {
            if (e->reversed) okvec_reverse(kmers);
            to_return = 1;
            return_val = e->mini;
            e->mini_hash = current_hash;
            e->mini_pos = 0;
            e->mini = (uint64_t)candidate_canon;
            e->reversed = (candidate_canon == e->rc_mini_candidate);
        }
        if (!e->reversed) {
            e->saved_kmer = okmer_make(e->current_kmer, e->mini_pos, m);
            e->saved_kmer.minimizer = e->mini;
        } else {
            e->saved_kmer = okmer_make(e->current_rc_kmer, (uint8_t)(k - m - e->mini_pos), m);
            e->saved_kmer.minimizer = e->mini;
        }
        e->saved_kmer.ridx = (uint32_t)e->seq_idx;
        e->saved_kmer.rev = (uint8_t)e->reversed;
        e->saved_kmer.pos = e->mini_pos;
        if (to_return && e->seq_idx > 0) {
            e->seq_idx += 1;
            e->saved = 1;
            return return_val;
        } else {
            if (e->seq_idx == 0) to_return = 0;
            okvec_push(kmers, e->saved_kmer);
        }
    }
    if (kmers->n > 0) {
        if (e->reversed) okvec_reverse(kmers);
        return e->mini;
    }
    return (kint)0;
}

/* kmer_full::replace_slice, brisk/Kmers.cpp:149-159 */
static void replace_slice(okmer *km, kint replacement, size_t position, size_t length) {
    kint mask = (((kint)1) << (2 * length)) - 1;
    replacement &= mask;
    mask <<= 2 * position;
    km->kmer_s &= ~mask;
    replacement <<= 2 * position;
    km->kmer_s += replacement;
}

/* kmer_full::hash_kmer_minimizer_inplace, brisk/Kmers.cpp:191-200 */
static void hash_kmer_minimizer_inplace(const oracle_params *p, okmer *km) {
    kint minimizer = km->kmer_s >> (km->minimizer_idx * 2);
    kint mask = (((kint)1) << (p->m * 2)) - 1;
    minimizer &= mask;
    minimizer = bfc_hash_64(p, (uint64_t)minimizer, (uint64_t)mask);
    km->minimizer = minimizer;
    replace_slice(km, minimizer, km->minimizer_idx, p->m);
}

/* kmer_full::unhash_kmer_minimizer, brisk/Kmers.cpp:178-187 */
static void unhash_kmer_minimizer(const oracle_params *p, okmer *km) {
    kint minimizer = km->kmer_s >> (km->minimizer_idx * 2);
    kint mask = (((kint)1) << (p->m * 2)) - 1;
    minimizer &= mask;
    minimizer = bfc_hash_64_inv((uint64_t)minimizer, (uint64_t)mask);
    km->minimizer = minimizer;
    replace_slice(km, minimizer, km->minimizer_idx, p->m);
}

/* kmer_full::get_compacted, brisk/Kmers.cpp:138-145 */
static kint get_compacted(const okmer *km, uint8_t m, uint8_t mini_idx) {
    kint mask = (((kint)1) << (mini_idx * 2)) - 1;
    kint suffix = km->kmer_s & mask;
    mask = ~mask;
    kint prefix = (km->kmer_s >> (2 * m));
    prefix = prefix & mask;
    return prefix + suffix;
}

/* bucket id: Brisk.hpp:135-137, DenseMenuYo.hpp:284-286 */
static uint64_t small_minimizer(const oracle_params *p, const okmer *hashed) {
    uint64_t sm = (uint64_t)hashed->minimizer >> (2 * ((p->m_reduc + 1) / 2));
    sm &= (uint64_t)((((kint)1) << (p->b * 2)) - 1);
    return sm;
}

/* ------------------------------------------------------------------- the store
 *
 * The reference keeps, per bucket id, a list of packed superkmers (SKL,
 * brisk/SuperKmerLight.hpp:17-24) and finds a k-mer by (position idx+s, compacted
 * k-b nucleotides) — SKL::is_kmer_present, SuperKmerLight.hpp:209-217 — inside the
 * bucket selected by the hashed minimizer's middle 2b bits.  (bucket id, compacted
 * key, position) is a bijection of (hashed kmer_s, minimizer_idx), so as a SET the
 * index is exactly { (hashed kmer_s, minimizer_idx) -> DATA }.  The oracle stores
 * that set in a plain open-addressing map; bucket ids are kept for stats().
 */
typedef struct {
    kint key;      /* kmer_s with the minimizer hashed in place */
    uint8_t idx;   /* minimizer_idx */
    uint8_t used;
    uint8_t data;  /* the counter app's uint8_t DATA */
} oslot;

typedef struct {
    oracle_params p;
    oslot *slots;
    uint64_t cap, n;
    okvec scratch;
} oracle_t;

static uint64_t okey_hash(kint key, uint8_t idx) {
    uint64_t x = (uint64_t)key ^ ((uint64_t)(key >> 64) * 0x9E3779B97F4A7C15ull) ^
                 ((uint64_t)idx << 57);
    x ^= x >> 31;
    x *= 0xBF58476D1CE4E5B9ull;
    x ^= x >> 29;
    x *= 0x94D049BB133111EBull;
    x ^= x >> 32;
    return x;
}

static void ostore_grow(oracle_t *o);

/* find-or-insert; returns the slot, *newly says whether it was created */
static oslot *ostore_upsert(oracle_t *o, kint key, uint8_t idx, int insert, int *newly) {
    if (insert && (o->n + 1) * 10 > o->cap * 6) ostore_grow(o);
    uint64_t mask = o->cap - 1;
    uint64_t i = okey_hash(key, idx) & mask;
    for (;;) {
        oslot *s = &o->slots[i];
        if (!s->used) {
            if (!insert) return NULL;
            s->used = 1;
            s->key = key;
            s->idx = idx;
            s->data = 0;
            o->n++;
            if (newly) *newly = 1;
            return s;
        }
        if (s->key == key && s->idx == idx) {
            if (newly) *newly = 0;
            return s;
        }
        i = (i + 1) & mask;
    }
}

static void ostore_grow(oracle_t *o) {
    oslot *old = o->slots;
    uint64_t oldcap = o->cap;
    o->cap = oldcap ? oldcap * 2 : (1u << 16);
    o->slots = (oslot *)calloc(o->cap, sizeof(oslot));
    o->n = 0;
    for (uint64_t i = 0; i < oldcap; i++) {
        if (old[i].used) {
            uint64_t mask = o->cap - 1;
            uint64_t j = okey_hash(old[i].key, old[i].idx) & mask;
            while (o->slots[j].used) j = (j + 1) & mask;
            o->slots[j] = old[i];
            o->n++;
        }
    }
    free(old);
}

/* ------------------------------------------------------------------ exported API */

void *oracle_new(int k, int m, int b) {
    if (!(m % 2 == 1) || !(m < k) || k > 64 || b > m || m > 31 || b < 1) return NULL; /* Brisk.hpp:50-51 */
    oracle_t *o = (oracle_t *)calloc(1, sizeof(oracle_t));
    o->p.k = k;
    o->p.m = m;
    o->p.b = b;
    o->p.m_reduc = m - b;
    o->p.s = (o->p.m_reduc + 1) / 2;
    o->p.k_mask = k >= 64 ? ~(kint)0 : (((kint)1 << (2 * k)) - 1); /* Q10 fix */
    o->p.m_mask = ((uint64_t)1 << (2 * m)) - 1;
    dec_init(&o->p);
    ostore_grow(o);
    return o;
}

void oracle_free(void *h) {
    oracle_t *o = (oracle_t *)h;
    if (!o) return;
    free(o->slots);
    free(o->scratch.v);
    free(o);
}

/* decycling class and full 64-bit order hash of one m-mer value */
unsigned oracle_mem_double(void *h, uint64_t mmer) { return dec_memDouble(&((oracle_t *)h)->p, mmer); }
uint64_t oracle_bfc_hash(void *h, uint64_t mmer) {
    oracle_t *o = (oracle_t *)h;
    return bfc_hash_64(&o->p, mmer, o->p.m_mask);
}
uint64_t oracle_bfc_hash_inv(void *h, uint64_t x) { return bfc_hash_64_inv(x, ((oracle_t *)h)->p.m_mask); }
uint64_t oracle_rcbc(uint64_t x, uint64_t n) { return rcbc(x, n); }
void oracle_rcb(uint64_t lo, uint64_t hi, uint64_t n, uint64_t *olo, uint64_t *ohi) {
    kint r = rcb(((kint)hi << 64) | lo, n);
    *olo = (uint64_t)r;
    *ohi = (uint64_t)(r >> 64);
}
double oracle_coef(void *h, int i) { return ((oracle_t *)h)->p.coef[i]; }

/* get_minimizer on an arbitrary 128-bit value (kk = k or k-1) */
uint64_t oracle_get_minimizer(void *h, uint64_t lo, uint64_t hi, int kk, uint8_t *pos, uint8_t *rev) {
    oracle_t *o = (oracle_t *)h;
    int r = 0;
    uint64_t mini = get_minimizer(&o->p, ((kint)hi << 64) | lo, (uint8_t)kk, pos, (uint8_t)o->p.m, &r, o->p.m_mask);
    *rev = (uint8_t)r;
    return mini;
}

/* the per-step candidate stream of SuperKmerEnumerator::next, brisk/Kmers.cpp:548-549:
 * for read k-mer i, hash = bfc_hash_64(min(mini_candidate, rc_mini_candidate)) and
 * rev = (canon == rc).  Returns the number of k-mers. */
int64_t oracle_hash_stream(void *h, const char *seq, uint64_t len, uint64_t *out_hash, uint8_t *out_rev) {
    oracle_t *o = (oracle_t *)h;
    int k = o->p.k, m = o->p.m;
    if (len < (uint64_t)k) return 0;
    kint mm = ((kint)1 << (2 * m)) - 1;
    kint f, r;
    init_kmer(seq + (k - m - 1), &f, &r, (uint8_t)m, mm >> 2);
    int64_t n = 0;
    for (uint64_t i = 0; k - 1 + i < len; i++) {
        update_kmer(seq[k - 1 + i], &f, &r, (uint8_t)m, mm);
        kint c = f < r ? f : r;
        out_hash[n] = bfc_hash_64(&o->p, (uint64_t)c, o->p.m_mask);
        out_rev[n] = (c == r);
        n++;
    }
    return n;
}

/* Runs the literal enumerator over one read.  K-mers come out in the order the
 * reference emits them (superkmer by superkmer, each already reversed when the
 * minimizer was on the reverse strand).  skmer_id[j] numbers the superkmers.
 * ridx[j] = index of the k-mer in the read, pos[j]/rev[j] = enumerator state. */
int64_t oracle_enumerate(void *h, const char *seq, uint64_t len, uint64_t cap, uint64_t *lo, uint64_t *hi,
                         uint8_t *idx, uint64_t *mini, uint32_t *skmer_id, uint32_t *ridx, uint8_t *pos,
                         uint8_t *rev) {
    oracle_t *o = (oracle_t *)h;
    if (len < (uint64_t)o->p.k) return 0; /* apps/counter.cpp:233 */
    oenum e;
    oenum_init(&e, &o->p, seq, len);
    okvec *v = &o->scratch;
    uint64_t n = 0;
    uint32_t sk = 0;
    v->n = 0;
    oenum_next(&e, v);
    while (v->n > 0) { /* apps/counter.cpp:243 */
        for (size_t j = 0; j < v->n; j++) {
            if (n >= cap) return -1;
            if (lo) lo[n] = (uint64_t)v->v[j].kmer_s;
            if (hi) hi[n] = (uint64_t)(v->v[j].kmer_s >> 64);
            if (idx) idx[n] = v->v[j].minimizer_idx;
            if (mini) mini[n] = (uint64_t)v->v[j].minimizer;
            if (skmer_id) skmer_id[n] = sk;
            if (ridx) ridx[n] = v->v[j].ridx;
            if (pos) pos[n] = v->v[j].pos;
            if (rev) rev[n] = v->v[j].rev;
            n++;
        }
        sk++;
        v->n = 0;
        oenum_next(&e, v);
    }
    return (int64_t)n;
}

/* insert_superkmer + the counter's DATA update, for the superkmer in o->scratch:
 * brisk/Brisk.hpp:123-147 (hash_skmer, bucket id from superkmer[0], lookup, insert
 * misses, lookup again, unhash) and apps/counter.cpp:262-269 (new -> 1, else ++). */
static void count_superkmer(oracle_t *o, okvec *v) {
    for (size_t j = 0; j < v->n; j++) {
        okmer km = v->v[j];
        hash_kmer_minimizer_inplace(&o->p, &km);
        int newly = 0;
        oslot *s = ostore_upsert(o, km.kmer_s, km.minimizer_idx, 1, &newly);
        if (newly)
            s->data = 1;
        else
            s->data++;
    }
}

/* count_sequence, apps/counter.cpp:231-276 */
void oracle_count_read(void *h, const char *seq, uint64_t len) {
    oracle_t *o = (oracle_t *)h;
    if (len < (uint64_t)o->p.k) return;
    oenum e;
    oenum_init(&e, &o->p, seq, len);
    okvec *v = &o->scratch;
    v->n = 0;
    oenum_next(&e, v);
    while (v->n > 0) {
        count_superkmer(o, v);
        v->n = 0;
        oenum_next(&e, v);
    }
}

void oracle_count_reads(void *h, const char *bases, const uint64_t *offsets, uint64_t n_reads) {
    for (uint64_t r = 0; r < n_reads; r++)
        oracle_count_read(h, bases + offsets[r], offsets[r + 1] - offsets[r]);
}

/* query_sequence, apps/counter.cpp:281-310, but reporting per read k-mer (read order):
 * found[i] / data[i] for k-mer i of the read; returns the sum of found counts
 * (the reference's `total`).  Note: the reference's loop stops early when a
 * superkmer's minimizer value is 0 (counter.cpp:304-306); `literal_break` = 1
 * reproduces that, 0 queries every superkmer. */
uint64_t oracle_query_read(void *h, const char *seq, uint64_t len, uint8_t *found, uint8_t *data, int literal_break) {
    oracle_t *o = (oracle_t *)h;
    if (len < (uint64_t)o->p.k) return 0;
    uint64_t total = 0;
    oenum e;
    oenum_init(&e, &o->p, seq, len);
    okvec *v = &o->scratch;
    v->n = 0;
    kint minimizer = oenum_next(&e, v);
    (void)minimizer;
    while (v->n > 0) {
        for (size_t j = 0; j < v->n; j++) {
            okmer km = v->v[j];
            hash_kmer_minimizer_inplace(&o->p, &km);
            oslot *s = ostore_upsert(o, km.kmer_s, km.minimizer_idx, 0, NULL);
            if (found) found[km.ridx] = s != NULL;
            if (data) data[km.ridx] = s ? s->data : 0;
            if (s) total += s->data;
        }
        v->n = 0;
        minimizer = oenum_next(&e, v);
        if (literal_break && minimizer == 0) break;
    }
    return total;
}

/* Brisk::insert of explicit (kmer_s, minimizer_idx) pairs — insert_superkmer on a
 * size-1 superkmer (Brisk.hpp:123-147).  newly[i] as in `newly_inserted`. */
void oracle_insert_kmers(void *h, const uint64_t *lo, const uint64_t *hi, const uint8_t *idx, uint64_t n,
                         uint8_t *newly, int count) {
    oracle_t *o = (oracle_t *)h;
    for (uint64_t i = 0; i < n; i++) {
        okmer km = okmer_make(((kint)hi[i] << 64) | lo[i], idx[i], (uint8_t)o->p.m);
        hash_kmer_minimizer_inplace(&o->p, &km);
        int nw = 0;
        oslot *s = ostore_upsert(o, km.kmer_s, km.minimizer_idx, 1, &nw);
        if (newly) newly[i] = (uint8_t)nw;
        if (count) { if (nw) s->data = 1; else s->data++; }
    }
}

/* Brisk::get, Brisk.hpp:64-69 */
void oracle_get_kmers(void *h, const uint64_t *lo, const uint64_t *hi, const uint8_t *idx, uint64_t n,
                      uint8_t *found, uint8_t *data) {
    oracle_t *o = (oracle_t *)h;
    for (uint64_t i = 0; i < n; i++) {
        okmer km = okmer_make(((kint)hi[i] << 64) | lo[i], idx[i], (uint8_t)o->p.m);
        hash_kmer_minimizer_inplace(&o->p, &km);
        oslot *s = ostore_upsert(o, km.kmer_s, km.minimizer_idx, 0, NULL);
        found[i] = s != NULL;
        data[i] = s ? s->data : 0;
    }
}

uint64_t oracle_size(void *h) { return ((oracle_t *)h)->n; }

typedef struct { kint key; uint8_t idx, data; } orec;
static int orec_cmp(const void *a, const void *b) {
    const orec *x = (const orec *)a, *y = (const orec *)b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return (int)x->idx - (int)y->idx;
}

/* The stored set as the reference exposes it through next()/get()
 * (Brisk.hpp:166-172 un-hashes the minimizer; apps/counter.cpp:94-104):
 * (kmer_s, minimizer_idx, DATA) sorted by (kmer_s, minimizer_idx). */
uint64_t oracle_dump(void *h, uint64_t *lo, uint64_t *hi, uint8_t *idx, uint8_t *data) {
    oracle_t *o = (oracle_t *)h;
    orec *r = (orec *)malloc((o->n + 1) * sizeof(orec));
    uint64_t n = 0;
    for (uint64_t i = 0; i < o->cap; i++) {
        if (!o->slots[i].used) continue;
        okmer km;
        km.kmer_s = o->slots[i].key;
        km.minimizer_idx = o->slots[i].idx;
        unhash_kmer_minimizer(&o->p, &km);
        r[n].key = km.kmer_s;
        r[n].idx = km.minimizer_idx;
        r[n].data = o->slots[i].data;
        n++;
    }
    qsort(r, n, sizeof(orec), orec_cmp);
    for (uint64_t i = 0; i < n; i++) {
        lo[i] = (uint64_t)r[i].key;
        hi[i] = (uint64_t)(r[i].key >> 64);
        idx[i] = r[i].idx;
        data[i] = r[i].data;
    }
    free(r);
    return n;
}

static int u64_cmp(const void *a, const void *b) {
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? -1 : x > y;
}

/* stats(), DenseMenuYo.hpp:545-568: buckets in use and stored k-mers.  (nb_skmers
 * and largest_bucket depend on insertion order in the reference and are not part
 * of the set; they are read from oracle/_ref when needed.) */
void oracle_stats(void *h, uint64_t *nb_buckets, uint64_t *nb_kmers) {
    oracle_t *o = (oracle_t *)h;
    uint64_t *ids = (uint64_t *)malloc((o->n + 1) * sizeof(uint64_t));
    uint64_t n = 0;
    for (uint64_t i = 0; i < o->cap; i++) {
        if (!o->slots[i].used) continue;
        okmer km;
        km.kmer_s = o->slots[i].key;
        km.minimizer_idx = o->slots[i].idx;
        km.minimizer = (km.kmer_s >> (2 * km.minimizer_idx)) & o->p.m_mask;
        ids[n++] = small_minimizer(&o->p, &km);
    }
    qsort(ids, n, sizeof(uint64_t), u64_cmp);
    uint64_t d = 0;
    for (uint64_t i = 0; i < n; i++)
        if (i == 0 || ids[i] != ids[i - 1]) d++;
    free(ids);
    *nb_buckets = d;
    *nb_kmers = o->n;
}

/* order hash of every m-mer of a read (offsets 0 .. len-m): bfc_hash_64(canonize(fwd))
 * as get_minimizer computes it per position (brisk/Kmers.cpp:372-373, 380-381).
 * rev[o] = 1 when the canonical form is the reverse complement. */
int64_t oracle_mmer_hashes(void *h, const char *seq, uint64_t len, uint64_t *out_hash, uint8_t *out_rev) {
    oracle_t *o = (oracle_t *)h;
    int m = o->p.m;
    if (len < (uint64_t)m) return 0;
    uint64_t f = 0;
    int64_t n = 0;
    for (uint64_t i = 0; i < len; i++) {
        f = ((f << 2) + nuc2int(seq[i])) & o->p.m_mask;
        if (i + 1 >= (uint64_t)m) {
            uint64_t c = canonize64(f, m);
            out_hash[n] = bfc_hash_64(&o->p, c, o->p.m_mask);
            if (out_rev) out_rev[n] = (c != f);
            n++;
        }
    }
    return n;
}
