/* TEST INFRASTRUCTURE ONLY - see shenqi_oracle.h.  CPU restatement of the reference's short-range
 * TreePM path in plain C.  Not part of the product; the product never calls into this file. */
#include "shenqi_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <omp.h>

#include "ph_tables.h"
#include "shortrange_table.h"

/* ------------------------------------------------------------------ Peano-Hilbert keys */
static const unsigned char ph_rot[48 * 8] = SQ_PH_ROT_INIT;
static const unsigned char ph_sub[48 * 8] = SQ_PH_SUB_INIT;

/* libgadget/utils/peano.cpp:144-183.  The reference walks two levels per step through a merged
 * table; one level per step gives the same key (the merged table is the composition, :119-139). */
uint64_t sqo_peano_hilbert_key(int x, int y, int z, int bits)
{
    unsigned rotation = 0;
    uint64_t key = 0;
    for(int bit = bits - 1; bit >= 0; bit--) {
        const unsigned pix = (((x >> bit) & 1) << 2) | (((y >> bit) & 1) << 1) | ((z >> bit) & 1);
        key = (key << 3) | ph_sub[rotation * 8 + pix];
        rotation = ph_rot[rotation * 8 + pix];
    }
    return key;
}

/* libgadget/utils/peano.h:15-21 */
void sqo_peano_keys(const double * pos, int64_t n, double box, uint64_t * keys)
{
    const double DomainFac = 1.0 / (box * 1.001) * (double) (((uint64_t) 1) << 21);
    #pragma omp parallel for
    for(int64_t i = 0; i < n; i++) {
        const double sx = pos[3 * i] + box / 2000, sy = pos[3 * i + 1] + box / 2000, sz = pos[3 * i + 2] + box / 2000;
        keys[i] = sqo_peano_hilbert_key((int) (sx * DomainFac), (int) (sy * DomainFac), (int) (sz * DomainFac), 21);
    }
}

/* libgadget/slotsmanager.cpp:395-437 */
struct sqo_order { uint64_t key; int64_t idx; unsigned char type; };
static int sqo_order_cmp(const void * a, const void * b)
{
    const struct sqo_order * pa = (const struct sqo_order *) a, * pb = (const struct sqo_order *) b;
    if(pa->type != pb->type) return pa->type < pb->type ? -1 : 1;
    if(pa->key != pb->key) return pa->key < pb->key ? -1 : 1;
    if(pa->idx != pb->idx) return pa->idx < pb->idx ? -1 : 1;
    return 0;
}
void sqo_sort_order(const uint64_t * keys, const unsigned char * types, int64_t n, int64_t * perm)
{
    struct sqo_order * o = (struct sqo_order *) malloc(sizeof(struct sqo_order) * (size_t) (n > 0 ? n : 1));
    for(int64_t i = 0; i < n; i++) {
        o[i].key = keys[i];
        o[i].idx = i;
        o[i].type = types ? types[i] : 1;
    }
    qsort(o, (size_t) n, sizeof(struct sqo_order), sqo_order_cmp);
    for(int64_t i = 0; i < n; i++)
        perm[i] = o[i].idx;
    free(o);
}

/* ------------------------------------------------------------------ force tree */
#define NMAXCHILD 8
#define NODEFULL (1 << 16)

static void set_childtype(sqo_node * nd, unsigned t) { nd->flags = (nd->flags & ~(3u << 3)) | (t << 3); }

/* libgadget/forcetree.cpp:277-283 */
static int get_subnode(const sqo_node * nd, const double * p)
{
    return (p[0] > nd->center[0]) + ((p[1] > nd->center[1]) << 1) + ((p[2] > nd->center[2]) << 2);
}

/* libgadget/forcetree.cpp:301-326 */
static void init_internal_node(sqo_node * nfreep, const sqo_node * parent, int subnode)
{
    const double lenhalf = 0.25 * parent->len;
    memset(nfreep, 0, sizeof(*nfreep));
    nfreep->len = 0.5 * parent->len;
    nfreep->sibling = -10;
    nfreep->father = -10;
    for(int j = 0; j < 3; j++) {
        const int sign = (subnode & (1 << j)) ? 1 : -1;
        nfreep->center[j] = parent->center[j] + sign * lenhalf;
    }
    for(int j = 0; j < NMAXCHILD; j++)
        nfreep->suns[j] = -1;
}

struct sqo_build { sqo_node * N; /* shifted: N[firstnode] is root */ int64_t firstnode, lastnode; int nnext;
                   const double * pos; const float * mass; };

/* libgadget/forcetree.cpp:946-952 (DM: no hmax) */
static void place_particle(struct sqo_build * b, int node, int slot, int p)
{
    sqo_node * nd = &b->N[node];
    nd->suns[slot] = p;
    nd->mass += b->mass[p];
    for(int k = 0; k < 3; k++)
        nd->cofm[k] += b->mass[p] * b->pos[3 * p + k];
}

/* libgadget/forcetree.cpp:369-476 */
static int create_new_node_layer(struct sqo_build * b, int firstparent, int p_toplace)
{
    int parent = firstparent;
    for(;;) {
        sqo_node * nprnt = &b->N[parent];
        int oldsuns[NMAXCHILD], newsuns[NMAXCHILD];
        memcpy(oldsuns, nprnt->suns, sizeof(oldsuns));
        if(b->nnext + 8 >= b->lastnode)
            return 1;
        newsuns[0] = b->nnext;
        b->nnext += 8;
        for(int i = 0; i < 8; i++) {
            newsuns[i] = newsuns[0] + i;
            init_internal_node(&b->N[newsuns[i]], nprnt, i);
            b->N[newsuns[i]].father = parent;
        }
        for(int i = 0; i < NMAXCHILD; i++) {
            const int sub = get_subnode(nprnt, b->pos + 3 * oldsuns[i]);
            sqo_node * nchild = &b->N[newsuns[sub]];
            place_particle(b, newsuns[sub], nchild->noccupied, oldsuns[i]);
            nchild->noccupied++;
        }
        memcpy(nprnt->suns, newsuns, sizeof(newsuns));
        for(int i = 0; i < 7; i++)
            b->N[newsuns[i]].sibling = newsuns[i + 1];
        b->N[newsuns[7]].sibling = nprnt->sibling;
        nprnt->mass = 0;
        nprnt->cofm[0] = nprnt->cofm[1] = nprnt->cofm[2] = 0;
        nprnt->hmax = 0;

        const int sub = get_subnode(nprnt, b->pos + 3 * p_toplace);
        const int child = nprnt->suns[sub];
        sqo_node * nchild = &b->N[child];
        if(nchild->noccupied < NMAXCHILD) {
            place_particle(b, child, nchild->noccupied, p_toplace);
            nchild->noccupied++;
            break;
        }
        set_childtype(nchild, SQO_NODE_NODE);
        nchild->noccupied = NODEFULL;
        parent = child;
    }
    set_childtype(&b->N[firstparent], SQO_NODE_NODE);
    b->N[firstparent].noccupied = NODEFULL;
    return 0;
}

/* libgadget/forcetree.cpp:480-519 */
static int add_particle_to_tree(struct sqo_build * b, int i, int cur)
{
    while(b->N[cur].noccupied >= NODEFULL)
        cur = b->N[cur].suns[get_subnode(&b->N[cur], b->pos + 3 * i)];
    const int nocc = b->N[cur].noccupied;
    b->N[cur].noccupied++;
    if(nocc < NMAXCHILD)
        place_particle(b, cur, nocc, i);
    else if(create_new_node_layer(b, cur, i))
        return -1;
    return cur;
}

/* libgadget/forcetree.cpp:984-1003 */
static void update_particle_node(sqo_node * nd)
{
    if(nd->mass > 0)
        for(int j = 0; j < 3; j++) nd->cofm[j] /= nd->mass;
    else
        for(int j = 0; j < 3; j++) nd->cofm[j] = nd->center[j];
}

/* libgadget/forcetree.cpp:1016-1103 */
static void update_node_recursive(struct sqo_build * b, int no, int sib)
{
    sqo_node * nd = &b->N[no];
    int * suns = nd->suns;
    int jj = 0;
    for(int j = 0; j < 8; j++, jj++) {
        while(jj < 8 && !SQO_TOPLEVEL(&b->N[suns[jj]]) && SQO_CHILDTYPE(&b->N[suns[jj]]) == SQO_PARTICLE_NODE
                && b->N[suns[jj]].noccupied == 0)
            jj++;
        suns[j] = jj < 8 ? suns[jj] : -1;
    }
    for(int j = 0; j < 8; j++) {
        const int p = suns[j];
        if(p < 0) continue;
        int nextsib = sib;
        for(int k = j + 1; k < 8; k++)
            if(suns[k] >= 0) { nextsib = suns[k]; break; }
        b->N[p].sibling = nextsib;
        if(SQO_CHILDTYPE(&b->N[p]) == SQO_PARTICLE_NODE)
            update_particle_node(&b->N[p]);
        else if(SQO_CHILDTYPE(&b->N[p]) == SQO_NODE_NODE)
            update_node_recursive(b, p, nextsib);
    }
    for(int j = 0; j < 8; j++) {
        const int p = suns[j];
        if(p < 0) continue;
        nd->mass += b->N[p].mass;
        for(int k = 0; k < 3; k++)
            nd->cofm[k] += b->N[p].mass * b->N[p].cofm[k];
        if(b->N[p].hmax > nd->hmax) nd->hmax = b->N[p].hmax;
    }
    if(nd->mass > 0)
        for(int k = 0; k < 3; k++) nd->cofm[k] /= nd->mass;
}

int64_t sqo_tree_build(const double * pos, const float * mass, int64_t n, double box,
                       sqo_node * nodes, int64_t maxnodes, int64_t firstnode)
{
    return sqo_tree_build_opt(pos, mass, n, box, nodes, maxnodes, firstnode, 1);
}

int64_t sqo_tree_build_opt(const double * pos, const float * mass, int64_t n, double box,
                           sqo_node * nodes, int64_t maxnodes, int64_t firstnode, int last_node_shortcut)
{
    struct sqo_build b;
    b.N = nodes - firstnode;
    b.firstnode = firstnode;
    b.lastnode = firstnode + maxnodes;
    b.nnext = (int) firstnode;
    b.pos = pos;
    b.mass = mass;
    /* libgadget/forcetree.cpp:653-686: root = 1.001 Box cube centred on Box/2, the single top leaf */
    sqo_node * root = &b.N[b.nnext++];
    memset(root, 0, sizeof(*root));
    root->len = box * 1.001;
    for(int i = 0; i < 3; i++) root->center[i] = box / 2.;
    for(int i = 0; i < NMAXCHILD; i++) root->suns[i] = -1;
    root->father = -1;
    root->sibling = -1;
    root->flags = 1u << 1; /* TopLevel */
    /* libgadget/forcetree.cpp:790-831, one thread */
    int this_acc = (int) firstnode;
    for(int64_t i = 0; i < n; i++) {
        const sqo_node * a = &b.N[this_acc];
        int cur = (int) firstnode;
        if(last_node_shortcut && fabs(2 * (pos[3 * i] - a->center[0])) <= a->len && fabs(2 * (pos[3 * i + 1] - a->center[1])) <= a->len
                && fabs(2 * (pos[3 * i + 2] - a->center[2])) <= a->len)
            cur = this_acc;
        this_acc = add_particle_to_tree(&b, (int) i, cur);
        if(this_acc < 0)
            return -1;
    }
    /* libgadget/forcetree.cpp:1118-1142 */
    root->flags |= 1u << 2; /* DependsOnLocalMass */
    if(SQO_CHILDTYPE(root) == SQO_NODE_NODE)
        update_node_recursive(&b, (int) firstnode, root->sibling);
    else
        update_particle_node(root);
    return b.nnext - firstnode;
}

/* ------------------------------------------------------------------ short-range window + walk */
static const double sr_x[SQ_NGRAVTAB] = SQ_SHORTRANGE_X_INIT;
static const double sr_pot[SQ_NGRAVTAB] = SQ_SHORTRANGE_POT_INIT;
static const double sr_force[SQ_NGRAVTAB] = SQ_SHORTRANGE_FORCE_INIT;

/* libgadget/gravshort-tree2.cpp:29-69 */
void sqo_window_table(int windowtype, double asmth, float * force_tab, float * pot_tab, double * dx)
{
    *dx = sr_x[1];
    for(int i = 0; i < SQ_NGRAVTAB; i++) {
        const double u = sr_x[i] * 0.5 / asmth;
        if(windowtype == 1) {
            force_tab[i] = (float) sr_force[i];
            pot_tab[i] = (float) sr_pot[i];
        } else {
            force_tab[i] = (float) (erfc(u) + 2.0 * u / sqrt(M_PI) * exp(-u * u));
            pot_tab[i] = (float) erfc(u);
        }
    }
}

struct sqo_walk {
    double BoxSize, cellsize, Rcut, G, h, ErrTolForceAcc, BHOpeningAngle2, dx;
    int TreeUseBH;
    float ftab[SQ_NGRAVTAB], ptab[SQ_NGRAVTAB];
};

/* libgadget/partmanager.h:99 */
static inline double nearest(double x, double box)
{
    return (x > 0.5 * box) ? (x - box) : ((x < -0.5 * box) ? (x + box) : x);
}

/* libgadget/gravshort2.hpp:21-55 */
static void sqo_walk_init(struct sqo_walk * w, const sqo_grav_params * par)
{
    w->BoxSize = par->BoxSize;
    w->cellsize = par->BoxSize / par->Nmesh;
    w->Rcut = par->TreeRcut * par->Asmth * w->cellsize;
    w->G = par->G;
    w->h = 2.8 * par->GravitySoftening;
    w->ErrTolForceAcc = par->ErrTolForceAcc;
    w->TreeUseBH = par->TreeUseBH;
    w->BHOpeningAngle2 = par->BHOpeningAngle * par->BHOpeningAngle;
    if(par->TreeUseBH == 0)
        w->BHOpeningAngle2 = par->MaxBHOpeningAngle * par->MaxBHOpeningAngle;
    sqo_window_table(par->WindowType, par->Asmth, w->ftab, w->ptab, &w->dx);
}

/* libgadget/gravshort2.hpp:326-358 with the table lookup of libgadget/gravity.h:48-60 */
static inline void apply_accn(const struct sqo_walk * w, double * acc, double * pot, const double dx[3], double r2, double mass)
{
    const double h = w->h;
    const double r = sqrt(r2);
    double fac = mass / (r2 * r);
    double facpot = -mass / r;
    if(r2 < h * h) {
        double wp;
        const double h3_inv = 1.0 / h / h / h;
        const double u = r / h;
        if(u < 0.5) {
            fac = mass * h3_inv * (10.666666666667 + u * u * (32.0 * u - 38.4));
            wp = -2.8 + u * u * (5.333333333333 + u * u * (6.4 * u - 9.6));
        } else {
            fac = mass * h3_inv * (21.333333333333 - 48.0 * u + 38.4 * u * u - 10.666666666667 * u * u * u - 0.066666666667 / (u * u * u));
            wp = -3.2 + 0.066666666667 / u + u * u * (10.666666666667 + u * (-16.0 + u * (9.6 - 2.133333333333 * u)));
        }
        facpot = mass / h * wp;
    }
    const double i = r / w->cellsize / w->dx;
    const size_t tabindex = (size_t) floor(i);
    if(tabindex >= SQ_NGRAVTAB - 1)
        return;
    fac *= (tabindex + 1 - i) * w->ftab[tabindex] + (i - tabindex) * w->ftab[tabindex + 1];
    facpot *= (tabindex + 1 - i) * w->ptab[tabindex] + (i - tabindex) * w->ptab[tabindex + 1];
    for(int k = 0; k < 3; k++)
        acc[k] += dx[k] * fac;
    *pot += facpot;
}

/* libgadget/gravshort2.hpp:152-167 */
static inline int discard_node(const sqo_node * nop, double r2, const double * inpos, const struct sqo_walk * w)
{
    if(r2 <= w->Rcut * w->Rcut)
        return 0;
    const double eff_dist = w->Rcut + 0.5 * nop->len;
    for(int i = 0; i < 3; i++)
        if(fabs(nearest(nop->center[i] - inpos[i], w->BoxSize)) > eff_dist)
            return 1;
    return 0;
}

/* libgadget/gravshort2.hpp:172-193 */
static inline int open_node(const sqo_node * nop, double r2, const double * inpos, double aold, const struct sqo_walk * w)
{
    const double len = nop->len;
    if((w->TreeUseBH == 0) && (nop->mass * len * len > r2 * r2 * aold))
        return 1;
    if(len * len / r2 > w->BHOpeningAngle2)
        return 1;
    const double inside = 0.6 * len;
    if(fabs(nearest(nop->center[0] - inpos[0], w->BoxSize)) < inside &&
       fabs(nearest(nop->center[1] - inpos[1], w->BoxSize)) < inside &&
       fabs(nearest(nop->center[2] - inpos[2], w->BoxSize)) < inside)
        return 1;
    return 0;
}

int64_t sqo_grav_short_tree(const sqo_node * nodes, int64_t firstnode, const double * pos, const float * mass,
                            const double * oldacc, const int * active, int64_t nactive,
                            const sqo_grav_params * par, double * accel, double * potential, int64_t * counts)
{
    struct sqo_walk w;
    sqo_walk_init(&w, par);
    const sqo_node * N = nodes - firstnode;
    int64_t total = 0;
    #pragma omp parallel for schedule(dynamic, 64) reduction(+: total)
    for(int64_t q = 0; q < nactive; q++) {
        const int64_t t = active ? active[q] : q;
        const double * inpos = pos + 3 * t;
        const double aold = w.ErrTolForceAcc * oldacc[t];
        double acc[3] = {0, 0, 0}, pot = 0;
        int64_t nint = 0;
        /* libgadget/gravshort2.hpp:227-322, TREEWALK_PRIMARY */
        int no = (int) firstnode;
        while(no >= 0) {
            const sqo_node * nop = &N[no];
            double dx[3];
            for(int i = 0; i < 3; i++)
                dx[i] = nearest(nop->cofm[i] - inpos[i], w.BoxSize);
            const double r2 = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2];
            if(discard_node(nop, r2, inpos, &w)) {
                no = nop->sibling;
                continue;
            }
            if(!open_node(nop, r2, inpos, aold, &w)) {
                no = nop->sibling;
                apply_accn(&w, acc, &pot, dx, r2, nop->mass);
                nint++;
                continue;
            }
            const unsigned ct = SQO_CHILDTYPE(nop);
            if(ct == SQO_PARTICLE_NODE) {
                for(int i = 0; i < nop->noccupied; i++) {
                    const int pp = nop->suns[i];
                    double d[3];
                    for(int j = 0; j < 3; j++)
                        d[j] = nearest(pos[3 * pp + j] - inpos[j], w.BoxSize);
                    apply_accn(&w, acc, &pot, d, d[0] * d[0] + d[1] * d[1] + d[2] * d[2], mass[pp]);
                    nint++;
                }
                no = nop->sibling;
            } else if(ct == SQO_PSEUDO_NODE)
                no = nop->sibling;
            else
                no = nop->suns[0];
        }
        /* libgadget/gravshort2.hpp:88-107, 136-146 */
        for(int k = 0; k < 3; k++)
            accel[3 * t + k] = acc[k] * w.G;
        if(potential && par->update_potential) {
            double p = pot;
            p += mass[t] / (w.h / 2.8);
            p -= 2.8372975 * pow(mass[t], 2.0 / 3) * pow(par->rho0, 1.0 / 3);
            potential[t] = p * w.G;
        }
        if(counts) counts[t] = nint;
        total += nint;
    }
    return total;
}

/* libgadget/gravshort-pair.c:22-127; neighbours are those with r <= Rcut (treewalk ngbiter, Hsml = Rcut) */
void sqo_grav_short_pair(const double * pos, const float * mass, int64_t n, const sqo_grav_params * par,
                         double * accel, double * potential)
{
    struct sqo_walk w;
    sqo_walk_init(&w, par);
    #pragma omp parallel for schedule(dynamic, 16)
    for(int64_t t = 0; t < n; t++) {
        double acc[3] = {0, 0, 0}, pot = 0;
        for(int64_t s = 0; s < n; s++) {
            double d[3];
            for(int j = 0; j < 3; j++)
                d[j] = nearest(pos[3 * s + j] - pos[3 * t + j], w.BoxSize);
            const double r2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
            if(r2 > w.Rcut * w.Rcut)
                continue;
            apply_accn(&w, acc, &pot, d, r2, mass[s]);
        }
        for(int k = 0; k < 3; k++)
            accel[3 * t + k] = acc[k] * w.G;
        if(potential) {
            double p = pot + mass[t] / (w.h / 2.8) - 2.8372975 * pow(mass[t], 2.0 / 3) * pow(par->rho0, 1.0 / 3);
            potential[t] = p * w.G;
        }
    }
}

/* libgadget/tests/test_gravity.cpp:44-83 (grav_force) + :133-153 (force_direct): Newtonian direct sum over the box and its
 * 26 mirror images with the spline softening below h = FORCE_SOFTENING(), unit masses scaled by mass[j], self pairs (r2 == 0)
 * skipped.  accn: N x 3, already x G. */
void sqo_direct_sum_27(const double * pos, const float * mass, int64_t n, double box, double G, double h, double * accn)
{
    const double h_inv = 1.0 / h, h3_inv = h_inv * h_inv * h_inv;
    #pragma omp parallel for schedule(dynamic, 16)
    for(int64_t i = 0; i < n; i++) {
        double a[3] = {0, 0, 0};
        for(int xx = -1; xx <= 1; xx++)
            for(int yy = -1; yy <= 1; yy++)
                for(int zz = -1; zz <= 1; zz++) {
                    const double off[3] = {box * xx, box * yy, box * zz};
                    for(int64_t j = 0; j < n; j++) {
                        double d[3], r2 = 0;
                        for(int k = 0; k < 3; k++) {
                            d[k] = off[k] + pos[3 * i + k] - pos[3 * j + k];
                            r2 += d[k] * d[k];
                        }
                        if(r2 == 0)
                            continue;
                        const double r = sqrt(r2);
                        double fac = 1 / (r2 * r);
                        if(r < h) {
                            const double u = r * h_inv;
                            if(u < 0.5)
                                fac = h3_inv * (10.666666666667 + u * u * (32.0 * u - 38.4));
                            else
                                fac = h3_inv * (21.333333333333 - 48.0 * u + 38.4 * u * u - 10.666666666667 * u * u * u - 0.066666666667 / (u * u * u));
                        }
                        for(int k = 0; k < 3; k++)
                            a[k] += -d[k] * fac * G * (mass ? mass[j] : 1.0);
                    }
                }
        for(int k = 0; k < 3; k++)
            accn[3 * i + k] = a[k];
    }
}

/* ------------------------------------------------------------------ particle mesh */
static inline int64_t wrapi(int64_t i, int n) { i %= n; return i < 0 ? i + n : i; }

/* libgadget/petapm.cpp:1132-1183 (stencil) + :1304-1310 (deposit); region = whole periodic mesh */
void sqo_cic_paint(const double * pos, const float * mass, int64_t n, double box, int nmesh, double * mesh)
{
    const double cell = box / nmesh;
    for(int64_t i = 0; i < n; i++) {
        int64_t ic[3]; double res[3];
        for(int k = 0; k < 3; k++) {
            const double tmp = pos[3 * i + k] / cell;
            ic[k] = (int64_t) floor(tmp);
            res[k] = tmp - ic[k];
        }
        for(int c = 0; c < 8; c++) {
            double weight = 1.0; int64_t lin = 0;
            for(int k = 0; k < 3; k++) {
                const int off = (c >> k) & 1;
                lin = lin * nmesh + wrapi(ic[k] + off, nmesh);
                weight *= off ? res[k] : (1 - res[k]);
            }
            mesh[lin] += weight * mass[i];
        }
    }
}

/* libgadget/petapm.cpp:1132-1183 + gravpm.cpp:489-500 */
void sqo_cic_readout(const double * mesh, const double * pos, int64_t n, double box, int nmesh, double * out)
{
    const double cell = box / nmesh;
    #pragma omp parallel for
    for(int64_t i = 0; i < n; i++) {
        int64_t ic[3]; double res[3];
        for(int k = 0; k < 3; k++) {
            const double tmp = pos[3 * i + k] / cell;
            ic[k] = (int64_t) floor(tmp);
            res[k] = tmp - ic[k];
        }
        for(int c = 0; c < 8; c++) {
            double weight = 1.0; int64_t lin = 0;
            for(int k = 0; k < 3; k++) {
                const int off = (c >> k) & 1;
                lin = lin * nmesh + wrapi(ic[k] + off, nmesh);
                weight *= off ? res[k] : (1 - res[k]);
            }
            out[i] += weight * mesh[lin];
        }
    }
}

/* libgadget/gravpm.cpp:295-302 */
static double sinc_unnormed(double x)
{
    if(x < 1e-5 && x > -1e-5) {
        const double x2 = x * x;
        return 1.0 - x2 / 6. + x2 * x2 / 120.;
    }
    return sin(x) / x;
}
/* libgadget/gravpm.cpp:436-444 */
static double diff_kernel(double w) { return 1 / 6.0 * (8 * sin(w) - sin(2 * w)); }

/* libgadget/gravpm.cpp:378-434 (potential_transfer without neutrinos/power spectrum) and :448-487 */
void sqo_pm_transfer(const double * rho_k, int nmesh, double box, double asmth, double G,
                     double * pot_k, double * fx_k, double * fy_k, double * fz_k)
{
    const int nz = nmesh / 2 + 1;
    const double asmth2 = pow((2 * M_PI) * asmth / nmesh, 2);
    const double pot_factor = -G / (M_PI * box);
    #pragma omp parallel for collapse(2)
    for(int ix = 0; ix < nmesh; ix++)
        for(int iy = 0; iy < nmesh; iy++)
            for(int iz = 0; iz < nz; iz++) {
                const int64_t ip = ((int64_t) ix * nmesh + iy) * nz + iz;
                /* libgadget/petapm.cpp:159-162 */
                int kpos[3] = {ix <= nmesh / 2 ? ix : ix - nmesh, iy <= nmesh / 2 ? iy : iy - nmesh, iz <= nmesh / 2 ? iz : iz - nmesh};
                const int64_t k2 = (int64_t) kpos[0] * kpos[0] + (int64_t) kpos[1] * kpos[1] + (int64_t) kpos[2] * kpos[2];
                double re = rho_k[2 * ip], im = rho_k[2 * ip + 1];
                if(k2 == 0) {
                    re = 0; im = 0;
                } else {
                    double f = 1.0;
                    const double smth = exp(-k2 * asmth2) / k2;
                    for(int k = 0; k < 3; k++) {
                        double tmp = (kpos[k] * M_PI) / nmesh;
                        tmp = sinc_unnormed(tmp);
                        f *= 1. / (tmp * tmp);
                    }
                    const double fac = pot_factor * smth * f * f;
                    re *= fac; im *= fac;
                }
                pot_k[2 * ip] = re; pot_k[2 * ip + 1] = im;
                double * outs[3] = {fx_k, fy_k, fz_k};
                for(int a = 0; a < 3; a++) {
                    const double fac = -1 * diff_kernel(kpos[a] * (2 * M_PI / nmesh)) * (nmesh / box);
                    outs[a][2 * ip] = -im * fac;
                    outs[a][2 * ip + 1] = re * fac;
                }
            }
}

/* ---------------------------------------------------------------------------------------------
 * drift, hierarchical gravity kick, active sub-list
 * ------------------------------------------------------------------------------------------- */

/* libgadget/drift.cpp:16-81.  Type 1 particles: no Hsml prediction (:54-67), no black-hole jump (:32-53). */
int64_t sqo_drift(double * pos, const double * vel, const unsigned char * skip, int64_t n, double ddrift,
                  const double * random_shift, double box)
{
    int64_t bad = 0;
    const double zero[3] = {0, 0, 0};
    const double * sh = random_shift ? random_shift : zero;
    #pragma omp parallel for reduction(+: bad)
    for(int64_t i = 0; i < n; i++) {
        double * p = pos + 3 * i;
        int ok = 1;
        if(skip && skip[i]) {            /* :20-29 */
            for(int j = 0; j < 3; j++)
                p[j] += sh[j];
        } else {
            for(int j = 0; j < 3; j++) { /* :68-75 */
                /* rounded product, no contraction; association (Pos + shift) + Vel*ddrift: what g++ -O3 -ffast-math
                 * (Options.mk.example:9) makes of drift.cpp:69, verified bit for bit against oracle/_ref; the
                 * source order Pos + (Vel*ddrift + shift) differs from it by at most 1 ulp. */
                volatile double t = vel[3 * i + j] * ddrift;
                volatile double q = p[j] + sh[j];
                p[j] = q + t;
                if(!isfinite(p[j])) { bad++; ok = 0; }
            }
        }
        if(!ok) continue;
        for(int j = 0; j < 3; j++) {     /* :76-79 */
            while(p[j] > box) p[j] -= box;
            while(p[j] <= 0) p[j] += box;
        }
    }
    return bad;
}

/* libgadget/timestep.cpp:270-279, 962-968 */
void sqo_kick(double * vel, const double * accel, const int32_t * active, const unsigned char * skip, int64_t nactive, double gravkick)
{
    #pragma omp parallel for
    for(int64_t i = 0; i < nactive; i++) {
        const int64_t pa = active ? active[i] : i;
        if(skip && skip[pa]) continue;
        for(int j = 0; j < 3; j++) {
            volatile double t = accel[3 * pa + j] * gravkick;
            vel[3 * pa + j] += t;
        }
    }
}

/* libgadget/timestep.cpp:132-139, dti_from_timebin libgadget/timebinmgr.h:42-45 */
int sqo_is_timebin_active(int bin, int64_t current)
{
    if(bin <= 0 || current <= 0) return 1;
    return current % (int64_t) (1ull << (uint64_t) bin) == 0;
}

/* libgadget/timestep.cpp:1353-1397 */
int64_t sqo_active_sublist(const int32_t * active, int64_t nactive, const unsigned char * timebin, const unsigned char * skip,
                           int maxtimebin, int64_t ti_current, int32_t * sublist)
{
    int64_t m = 0;
    for(int64_t i = 0; i < nactive; i++) {
        const int32_t pi = active ? active[i] : (int32_t) i;
        const int bin = timebin[pi];
        if(skip && skip[pi]) continue;
        if(bin > maxtimebin) continue;
        if(!sqo_is_timebin_active(bin, ti_current)) continue;
        sublist[m++] = pi;
    }
    return m;
}

/* libgadget/timebinmgr.cpp:189-202 */
int64_t sqo_round_down_power_of_two(int64_t dti)
{
    int64_t ti_min = (int64_t) 1 << 46;
    int sign = 1;
    if(dti < 0) { dti = -dti; sign = -1; }
    while(ti_min > dti) ti_min >>= 1;
    return ti_min * sign;
}

/* libgadget/timestep.cpp:1236-1250 */
int sqo_get_timestep_bin(int64_t dti)
{
    int bin = -1;
    if(dti <= 1) return 0;
    while(dti) { bin++; dti >>= 1; }
    return bin;
}

/* libgadget/timestep.cpp:351-377; volatile temporaries keep every product / sum separately rounded in source order. */
int64_t sqo_gravity_timebins(const sqo_timebin_params * P, const double * accel, const double * gravpm, const int32_t * active,
                             const unsigned char * skip, int64_t nactive, unsigned char * timebin, int64_t * counts)
{
    int64_t nbad = 0;
    for(int k = 0; k <= 46; k++) counts[k] = 0;
    for(int64_t i = 0; i < nactive; i++) {
        const int64_t pa = active ? active[i] : i;
        if(skip && skip[pa]) continue;
        const double * ga = accel + 3 * pa, * pm = gravpm + 3 * pa;
        /* grav_acceleration2, :1012-1027 */
        volatile double aa = P->atime * P->atime;
        volatile double a2inv = 1 / aa;
        volatile double a[3];
        for(int j = 0; j < 3; j++) {
            volatile double t1 = a2inv * ga[j], t2 = a2inv * pm[j];
            a[j] = t1 + t2;
        }
        volatile double s0 = a[0] * a[0], s1 = a[1] * a[1], s2 = a[2] * a[2];
        volatile double s01 = s0 + s1;
        volatile double ac2 = s01 + s2;
        if(ac2 == 0) ac2 = 1.0e-60;
        /* get_timestep_gravity_dloga, :1029-1039 */
        volatile double ac = sqrt(ac2);
        volatile double n1 = 2 * P->ErrTolIntAccuracy;
        volatile double n2 = n1 * P->atime;
        volatile double n3 = n2 * P->softening;
        volatile double q = n3 / ac;
        volatile double dt = sqrt(q);
        volatile double dloga = dt * P->hubble;
        /* convert_timestep_to_ti, :157-174 */
        int64_t dti;
        if(P->dti_max == 0)
            dti = 0;
        else {
            if(dloga < P->MinSizeTimestep) dloga = P->MinSizeTimestep;
            /* dti_from_dloga, timebinmgr.h:135-156; ti_from_loga_snap :123-132 */
            volatile double target = dloga + P->loga;
            int64_t snap = P->lastsnap; int k = 0;
            if(P->can_advance && P->sync_loga[1] <= target) { snap++; k = 1; }
            volatile double seg = P->sync_loga[k + 1] - P->sync_loga[k];
            volatile double logDTime = seg / (double) ((uint64_t) 1 << 46);
            int64_t tip = snap << 46;
            volatile double off = target - P->sync_loga[k];
            volatile double frac = off / logDTime;
            volatile double tnew = (double) tip + frac;
            tip = (tnew >= 9223372036854775808.0 || tnew < -9223372036854775808.0 || tnew != tnew) ? INT64_MIN : (int64_t) tnew;
            dti = tip - P->Ti_Current;
            if(dti > P->dti_max || dti < 0) dti = P->dti_max;
        }
        dti = sqo_round_down_power_of_two(dti);
        if(dti <= 1 || dti > ((int64_t) 1 << 46)) nbad++;
        int bin = sqo_get_timestep_bin(dti);
        if(bin > P->largest_active) bin = P->largest_active;
        counts[bin]++;
        timebin[pa] = (unsigned char) bin;
    }
    return nbad;
}

/* libgadget/timestep.cpp:379-398 */
int sqo_pushdown_bin(int64_t * counts, int largest_active, int isPM)
{
    for(int ti = largest_active; ti >= 1; ti--)
        if(counts[ti] > 0) { largest_active = ti; break; }
    int push_down_bin = largest_active;
    if(isPM) {
        for(int ti = largest_active; ti >= 1; ti--) {
            if(counts[ti] / 3 > counts[ti - 1]) break;
            push_down_bin = ti - 1;
            counts[ti - 1] += counts[ti];
        }
    }
    return push_down_bin;
}
