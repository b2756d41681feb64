/* mrb_feature_mc.h -- per-cell arithmetic of the feature-preserving marching cubes (--feature-detection),
 * shared by the CUDA kernels (mrb_feature_mc.cu) and the CPU oracle.
 *
 * Restates isFeature / getTransform / getVec / getNormMat / computePoint (marching_cubes/mcubes.cpp:9-152), the
 * feature branch of Polygonise (mcubes.cpp:234-311), getEdges / sameEdge (mcubes.cpp:396-425) and the per-pair body
 * of flipEdges (mcubes.cpp:434-588).  The least-squares feature point goes through include/mrb_eig3.h instead of
 * Eigen::JacobiSVD, so this sub-path is PARITY UNPINNED against the reference; sharing one definition
 * between oracle and GPU makes GPU == oracle bit for bit (fp32, no contraction on either side).  The loops AROUND
 * these functions (cell order, neighbour order, triangle concatenation) are written separately in the oracle and in
 * the kernels.
 */
#ifndef MRB_FEATURE_MC_H
#define MRB_FEATURE_MC_H

#include <stdint.h>
#include "mrb_eig3.h"

#define MRB_FEAT_MAX_TRI 15     /* 5 table triangles -> at most 15 boundary edges -> 15 fan triangles */

typedef struct MrbP3 { float x, y, z; } MrbP3;
typedef struct MrbTri { MrbP3 p[3]; } MrbTri;
typedef struct MrbFeatCell {
    int feature;                /* 0 none, 1 edge, 2 corner (mcubes.h:305-308) */
    int ntri;
    MrbP3 fp;                   /* feature point (-1,-1,-1 when feature == 0) */
    MrbTri tri[MRB_FEAT_MAX_TRI];
} MrbFeatCell;

MRB_HD int mrb_p3_less(MrbP3 a, MrbP3 b) {            /* XYZ::operator<, hasher.h:41-58 */
    if (a.x < b.x) return 1; else if (a.x > b.x) return 0;
    if (a.y < b.y) return 1; else if (a.y > b.y) return 0;
    if (a.z < b.z) return 1;
    return 0;
}
MRB_HD float mrb_p3_dist(MrbP3 l, MrbP3 r) {          /* XYZ::dist, hasher.h:28-33 */
    float d = (r.x - l.x) * (r.x - l.x);
    d += (r.y - l.y) * (r.y - l.y);
    d += (r.z - l.z) * (r.z - l.z);
    return sqrtf(d);
}
MRB_HD int mrb_p3_eq(MrbP3 l, MrbP3 r) { return (double)mrb_p3_dist(l, r) < 0.000001; }   /* XYZ::operator== */
MRB_HD int mrb_tri_degenerate(const MrbTri* t) {      /* mcubes.cpp:227-229 */
    if ((double)mrb_p3_dist(t->p[0], t->p[1]) < 0.00001) return 1;
    if ((double)mrb_p3_dist(t->p[1], t->p[2]) < 0.00001) return 1;
    if ((double)mrb_p3_dist(t->p[2], t->p[0]) < 0.00001) return 1;
    return 0;
}
MRB_HD MrbP3 mrb_vertex_interp(float iso, MrbP3 p1, MrbP3 p2, float v1, float v2) {   /* mcubes.cpp:369-392 */
    MrbP3 p;
    if (mrb_p3_less(p2, p1)) { MrbP3 t = p1; float tv = v1; p1 = p2; p2 = t; v1 = v2; v2 = tv; }
    {
        float dv = v1 - v2;
        if ((double)(dv < 0.0f ? -dv : dv) > 0.00001) {
            p.x = p1.x + (p2.x - p1.x) / (v2 - v1) * (iso - v1);
            p.y = p1.y + (p2.y - p1.y) / (v2 - v1) * (iso - v1);
            p.z = p1.z + (p2.z - p1.z) / (v2 - v1) * (iso - v1);
        } else {
            p = p1;
        }
    }
    return p;
}
MRB_HD int mrb_same_edge(MrbP3 a0, MrbP3 a1, MrbP3 b0, MrbP3 b1) {   /* sameEdge, mcubes.cpp:409-425 */
    if (mrb_p3_less(a1, a0)) { MrbP3 t = a0; a0 = a1; a1 = t; }
    if (mrb_p3_less(b1, b0)) { MrbP3 t = b0; b0 = b1; b1 = t; }
    return ((double)mrb_p3_dist(a0, b0) < 0.00001) && ((double)mrb_p3_dist(a1, b1) < 0.00001);
}

/* isFeature, mcubes.cpp:9-54.  nidx[8]: normal index per corner or -1. */
MRB_HD int mrb_is_feature(const int nidx[8], const float* normals, float feat_thresh, float corner_thresh) {
    float n0[3] = {0, 0, 0}, n1[3] = {0, 0, 0}, ns[3];
    float mn = 5.0f, max_ang = 0.0f;
    int n, m;
    for (n = 0; n < 7; n++) {
        if (nidx[n] == -1) continue;
        for (m = n + 1; m < 8; m++) {
            const float *a, *b;
            float angle;
            if (nidx[m] == -1) continue;
            a = normals + 3 * nidx[n]; b = normals + 3 * nidx[m];
            angle = a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
            if (angle < mn) { mn = angle; n0[0] = a[0]; n0[1] = a[1]; n0[2] = a[2]; n1[0] = b[0]; n1[1] = b[1]; n1[2] = b[2]; }
        }
    }
    if (mn >= feat_thresh) return 0;
    ns[0] = n0[1] * n1[2] - n0[2] * n1[1];
    ns[1] = n0[2] * n1[0] - n0[0] * n1[2];
    ns[2] = n0[0] * n1[1] - n0[1] * n1[0];
    for (n = 0; n < 8; n++) {
        const float* a;
        float c;
        if (nidx[n] == -1) continue;
        a = normals + 3 * nidx[n];
        c = a[0] * ns[0] + a[1] * ns[1] + a[2] * ns[2];
        if (c < 0.0f) c = -c;
        if (c > max_ang) max_ang = c;
    }
    return max_ang > corner_thresh ? 2 : 1;
}

/* Corner n of cell (i,j,k): (i + [n%4 in {1,2}], j + (n%4)/2, k + n/4), mcubes.cpp:166-168. */
MRB_HD MrbP3 mrb_cell_corner(int i, int j, int k, int n) {
    MrbP3 p;
    p.x = (float)(i + ((((n % 4) == 1) || ((n % 4) == 2)) ? 1 : 0));
    p.y = (float)(j + (n % 4) / 2);
    p.z = (float)(k + n / 4);
    return p;
}

/* Table triangles of the case word `w` (bits 0..3 count, nibbles = edge ids) with the degenerate filter. */
MRB_HD int mrb_table_triangles(uint64_t w, const MrbP3 vert[12], MrbTri* out) {
    int ntri = (int)(w & 15), t, n = 0;
    for (t = 0; t < ntri; t++) {
        MrbTri tr;
        tr.p[0] = vert[(w >> (4 + 12 * t)) & 15];
        tr.p[1] = vert[(w >> (8 + 12 * t)) & 15];
        tr.p[2] = vert[(w >> (12 + 12 * t)) & 15];
        if (mrb_tri_degenerate(&tr)) continue;
        out[n++] = tr;
    }
    return n;
}

/* Polygonise (mcubes.cpp:194-312) for one surface-crossing cell with USING_FEATURES.
 * val / nidx: the 8 corner values and normal indexes; w: case word of the cube index; ea/eb: edge -> corner tables. */
MRB_HD void mrb_polygonise_feature(int i, int j, int k, const float val[8], const int nidx[8], const float* normals,
                                   float iso, float feat_thresh, float corner_thresh, uint64_t w, int cubeindex,
                                   const int ea[12], const int eb[12], MrbFeatCell* cell) {
    MrbP3 corner[8], vert[12];
    int valid[12];
    int e, n, nv = 0;
    for (n = 0; n < 8; n++) corner[n] = mrb_cell_corner(i, j, k, n);
    for (e = 0; e < 12; e++) {                                   /* getSamplePoints, mcubes.cpp:314-367 */
        valid[e] = (((cubeindex >> ea[e]) & 1) != ((cubeindex >> eb[e]) & 1));
        if (valid[e]) { vert[e] = mrb_vertex_interp(iso, corner[ea[e]], corner[eb[e]], val[ea[e]], val[eb[e]]); nv++; }
        else { vert[e].x = -1.0f; vert[e].y = -1.0f; vert[e].z = -1.0f; }
    }
    cell->feature = mrb_is_feature(nidx, normals, feat_thresh, corner_thresh);   /* getGridCells, mcubes.cpp:176-183 */
    cell->fp.x = cell->fp.y = cell->fp.z = -1.0f;
    cell->ntri = 0;
    if (cell->feature != 0) {
        /* getTransform (:57-74), getVec (:76-129), getNormMat (:131-138), computePoint (:141-152) */
        float shift[3] = {0, 0, 0}, M[36], vec[12], fpnt[3];
        int rows = 0, ok = 1;
        for (e = 0; e < 12; e++)
            if (valid[e]) { shift[0] += vert[e].x; shift[1] += vert[e].y; shift[2] += vert[e].z; }
        shift[0] /= (float)nv; shift[1] /= (float)nv; shift[2] /= (float)nv;
        for (e = 0; e < 12 && ok; e++) {
            float pd[8];
            int pn[8], m, index = -1;
            const float* nm;
            if (!valid[e]) continue;
            for (n = 0; n < 8; n++) {
                float dist = 0.0f;
                dist += (corner[n].x - vert[e].x) * (corner[n].x - vert[e].x);
                dist += (corner[n].y - vert[e].y) * (corner[n].y - vert[e].y);
                dist += (corner[n].z - vert[e].z) * (corner[n].z - vert[e].z);
                pd[n] = dist; pn[n] = n;
            }
            for (n = 1; n < 8; n++)                              /* the reference's bubble sort, min to max */
                for (m = 0; m < 8 - n; m++)
                    if (pd[m] > pd[m + 1]) { float td = pd[m]; int ti = pn[m]; pd[m] = pd[m + 1]; pn[m] = pn[m + 1]; pd[m + 1] = td; pn[m + 1] = ti; }
            for (n = 0; n < 8; n++) { if (nidx[pn[n]] == -1) continue; index = pn[n]; break; }
            if (index < 0) { ok = 0; break; }                    /* reference: normals[normal_indexes[-1]] (UB); here: no feature */
            nm = normals + 3 * nidx[index];
            vec[rows] = (vert[e].x - shift[0]) * nm[0] + (vert[e].y - shift[1]) * nm[1] + (vert[e].z - shift[2]) * nm[2];
            M[rows * 3 + 0] = nm[0]; M[rows * 3 + 1] = nm[1]; M[rows * 3 + 2] = nm[2];
            rows++;
        }
        if (!ok) {
            cell->feature = 0;
        } else {
            float xd, yd, zd, dist;
            mrb_feature_point(M, vec, rows, shift, cell->feature == 1, fpnt);
            cell->fp.x = fpnt[0]; cell->fp.y = fpnt[1]; cell->fp.z = fpnt[2];
            xd = (cell->fp.x - shift[0]) * (cell->fp.x - shift[0]);
            yd = (cell->fp.y - shift[1]) * (cell->fp.y - shift[1]);
            zd = (cell->fp.z - shift[2]) * (cell->fp.z - shift[2]);
            dist = sqrtf(xd + yd + zd);
            if (dist < 1.0f) {                                   /* mcubes.cpp:249-293: fan over the boundary edges */
                MrbTri tmp[5];
                int nt = mrb_table_triangles(w, vert, tmp), a, b, ne = nt * 3;
                for (a = 0; a < ne; a++) {
                    MrbP3 a0 = tmp[a / 3].p[a % 3], a1 = tmp[a / 3].p[(a % 3 + 1) % 3];
                    int add = 1;
                    MrbTri t;
                    for (b = 0; b < ne; b++) {
                        if (a == b) continue;
                        if (mrb_same_edge(a0, a1, tmp[b / 3].p[b % 3], tmp[b / 3].p[(b % 3 + 1) % 3])) add = 0;
                    }
                    if (!add) continue;
                    t.p[0] = a0; t.p[1] = a1; t.p[2] = cell->fp;
                    if (mrb_tri_degenerate(&t)) continue;
                    if (cell->ntri < MRB_FEAT_MAX_TRI) cell->tri[cell->ntri++] = t;
                }
                return;
            }
            cell->feature = 0;                                   /* feature point too far: plain triangles (:296-309) */
        }
    }
    cell->ntri = mrb_table_triangles(w, vert, cell->tri);
}

/* Body of flipEdges for one (cell, neighbour) pair of feature cells (mcubes.cpp:493-582): finds the triangle pairs
 * that share an edge, emits the two flipped triangles per pair into out[] (returns how many) and removes the pair
 * from both cells. */
/* do triangle i of cell c and triangle j of cell nb share an edge? (mcubes.cpp:507-523) */
MRB_HD int mrb_flip_shared(const MrbFeatCell* c, const MrbFeatCell* nb, int i, int j) {
    int n, m;
    for (n = 0; n < 3; n++)
        for (m = 0; m < 3; m++)
            if (mrb_same_edge(c->tri[i].p[n], c->tri[i].p[(n + 1) % 3], nb->tri[j].p[m], nb->tri[j].p[(m + 1) % 3])) return 1;
    return 0;
}
/* the sequential part (mcubes.cpp:527-582): si/sj = the sharing (i,j) pairs in (i,j) order */
MRB_HD int mrb_flip_apply(MrbFeatCell* c, MrbFeatCell* nb, int* si, int* sj, int ns, MrbTri* out, int out_cap) {
    int i, t, nout = 0;
    for (t = 0; t < ns; t++) {
        MrbTri t1 = c->tri[si[t]], t2 = nb->tri[sj[t]];
        for (i = 0; i < 3; i++) if (mrb_p3_eq(t1.p[i], c->fp)) t1.p[(i + 2) % 3] = nb->fp;
        for (i = 0; i < 3; i++) if (mrb_p3_eq(t2.p[i], nb->fp)) t2.p[(i + 2) % 3] = c->fp;
        if (!mrb_tri_degenerate(&t1) && nout < out_cap) out[nout++] = t1;
        if (!mrb_tri_degenerate(&t2) && nout < out_cap) out[nout++] = t2;
        for (i = si[t]; i < c->ntri - 1; i++) c->tri[i] = c->tri[i + 1];
        c->ntri--;
        for (i = sj[t]; i < nb->ntri - 1; i++) nb->tri[i] = nb->tri[i + 1];
        nb->ntri--;
        for (i = t + 1; i < ns; i++) {
            if (si[i] > si[t]) si[i] -= 1;
            if (sj[i] > sj[t]) sj[i] -= 1;
        }
    }
    return nout;
}
MRB_HD int mrb_flip_pair(MrbFeatCell* c, MrbFeatCell* nb, MrbTri* out, int out_cap) {
    int si[MRB_FEAT_MAX_TRI * MRB_FEAT_MAX_TRI], sj[MRB_FEAT_MAX_TRI * MRB_FEAT_MAX_TRI];
    int ns = 0, i, j;
    for (i = 0; i < c->ntri; i++)
        for (j = 0; j < nb->ntri; j++)
            if (mrb_flip_shared(c, nb, i, j)) { si[ns] = i; sj[ns] = j; ns++; }
    return mrb_flip_apply(c, nb, si, sj, ns, out, out_cap);
}

#endif /* MRB_FEATURE_MC_H */
