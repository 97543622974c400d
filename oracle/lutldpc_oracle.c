/* TEST INFRASTRUCTURE ONLY -- see lutldpc_oracle.h.  Plain C99 restatement of the reference's
 * LUT-LDPC decode path; scalar, one frame at a time, int arithmetic like the reference. */
#include "lutldpc_oracle.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

enum { NODE_IM = 0, NODE_ROOT = 1, NODE_MSG = 2, NODE_CHA = 3 }; /* src/LUT_Tree.hpp:188-194 */

typedef struct {
  int type, K, nchild, qlen;
  int* Q;
  int* child; /* indices into the tree's node array */
} orc_node;

typedef struct {
  int tree_type, num_leaves, n_nodes;
  orc_node* nodes; /* pre-order, nodes[0] = root */
} orc_tree;

struct orc_decoder {
  int nvar, nchk, nchk_lin_indep, num_edges, max_iters, Nq_Cha, minLUT, n_qb_msg, has_gen;
  int psc, pisc;
  int32_t *dv, *dc, *cn_msg_idx, *Nq_Msg, *reuse, *cha2msg;
  int32_t* chk_equ_idx; /* flat, row by row: VN index of every entry of cn_msg_idx */
  double *qb_Cha, *qb_Msg;
  int n_var_iter, n_var_deg, n_chk_iter, n_chk_deg;
  orc_tree* var_trees; /* [n_var_iter][n_var_deg] */
  orc_tree* chk_trees;
  int32_t *var_tree_idx_iter, *var_tree_idx_degree, *chk_tree_idx_degree;
  int gen_rows, gen_cols;
  uint8_t* gen; /* [gen_rows][gen_cols] */
  int32_t* msgs; /* scratch, like the member LDPC_Code_LUT::msgs */
};

/* ---------- text cursor over the reference's tree serialisation ---------- */
typedef struct { const char* p; const char* end; } cursor;

static int next_int(cursor* c, int* out) {
  while (c->p < c->end && (*c->p == ' ' || *c->p == '\n' || *c->p == '\r' || *c->p == '\t')) c->p++;
  if (c->p >= c->end) return 0;
  int sign = 1, v = 0, nd = 0;
  if (*c->p == '-') { sign = -1; c->p++; }
  while (c->p < c->end && *c->p >= '0' && *c->p <= '9') { v = v * 10 + (*c->p - '0'); c->p++; nd++; }
  if (!nd) return 0;
  *out = sign * v;
  return 1;
}

/* Node record, src/LUT_Tree.cpp:507-517 (serialize_tree) / :448-535 (deserialise):
 *   <#children> \n <type> <|Q|> <K> \n [Q...] \n  then the children, pre-order. */
static int parse_node(cursor* c, orc_tree* t, int* next_free, int cap) {
  int nchild, type, qlen, K;
  if (!next_int(c, &nchild) || !next_int(c, &type) || !next_int(c, &qlen) || !next_int(c, &K)) return -1;
  if (*next_free >= cap) return -1;
  int me = (*next_free)++;
  orc_node* nd = &t->nodes[me];
  nd->type = type; nd->K = K; nd->nchild = nchild; nd->qlen = qlen;
  nd->Q = qlen ? (int*)malloc(sizeof(int) * (size_t)qlen) : NULL;
  nd->child = nchild ? (int*)malloc(sizeof(int) * (size_t)nchild) : NULL;
  for (int i = 0; i < qlen; i++) if (!next_int(c, &nd->Q[i])) return -1;
  for (int i = 0; i < nchild; i++) {
    int ch = parse_node(c, t, next_free, cap);
    if (ch < 0) return -1;
    t->nodes[me].child[i] = ch;
  }
  return me;
}

/* Array<Array<LUT_Tree>>, src/LUT_Tree.cpp:855-864: <#iter> { <#deg> { <tree_type> <num_leaves> tree } } */
static orc_tree* parse_tree_array(const char* txt, int len, int* n_iter, int* n_deg) {
  cursor c = { txt, txt + len };
  *n_iter = 0; *n_deg = 0;
  int ni;
  if (!next_int(&c, &ni) || ni <= 0) return NULL;
  orc_tree* arr = NULL;
  for (int ii = 0; ii < ni; ii++) {
    int nd;
    if (!next_int(&c, &nd)) return NULL;
    if (ii == 0) { *n_deg = nd; arr = (orc_tree*)calloc((size_t)ni * nd, sizeof(orc_tree)); }
    else if (nd != *n_deg) return NULL;
    for (int dd = 0; dd < nd; dd++) {
      orc_tree* t = &arr[ii * nd + dd];
      if (!next_int(&c, &t->tree_type) || !next_int(&c, &t->num_leaves)) return NULL;
      int cap = 2 * t->num_leaves + 2; /* every inner node has >= 1 child; generous for chains */
      if (cap < 8) cap = 8;
      cap *= 2;
      t->nodes = (orc_node*)calloc((size_t)cap, sizeof(orc_node));
      int nf = 0;
      if (parse_node(&c, t, &nf, cap) != 0) return NULL;
      t->n_nodes = nf;
    }
  }
  *n_iter = ni;
  return arr;
}

static int rd(FILE* f, void* dst, size_t n) { return fread(dst, 1, n, f) == n; }
static int32_t* rd_i32(FILE* f, int n) {
  int32_t* p = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
  if (n > 0 && !rd(f, p, sizeof(int32_t) * (size_t)n)) { free(p); return NULL; }
  return p;
}
static double* rd_f64(FILE* f, int n) {
  double* p = (double*)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
  if (n > 0 && !rd(f, p, sizeof(double) * (size_t)n)) { free(p); return NULL; }
  return p;
}
#define FAIL(msg) do { if (err) snprintf(err, (size_t)errlen, "%s", msg); if (f) fclose(f); return NULL; } while (0)

orc_decoder* orc_load_bundle(const char* path, char* err, int errlen) {
  FILE* f = fopen(path, "rb");
  if (!f) FAIL("cannot open bundle");
  char magic[8];
  int32_t h[10];
  if (!rd(f, magic, 8) || memcmp(magic, "LUTLDPC1", 8) != 0) FAIL("bad magic");
  if (!rd(f, h, sizeof h)) FAIL("short header");
  orc_decoder* d = (orc_decoder*)calloc(1, sizeof *d);
  d->nvar = h[0]; d->nchk = h[1]; d->nchk_lin_indep = h[2]; d->num_edges = h[3];
  d->max_iters = h[4]; d->Nq_Cha = h[5]; d->minLUT = h[6]; d->n_qb_msg = h[7]; d->has_gen = h[8];
  d->psc = 1; d->pisc = 1;
  d->dv = rd_i32(f, d->nvar); d->dc = rd_i32(f, d->nchk); d->cn_msg_idx = rd_i32(f, d->num_edges);
  d->Nq_Msg = rd_i32(f, d->max_iters); d->reuse = rd_i32(f, d->max_iters);
  d->cha2msg = rd_i32(f, d->Nq_Cha);
  d->qb_Cha = rd_f64(f, d->Nq_Cha - 1); d->qb_Msg = rd_f64(f, d->n_qb_msg);
  if (!d->dv || !d->dc || !d->cn_msg_idx || !d->Nq_Msg || !d->reuse || !d->cha2msg || !d->qb_Cha || !d->qb_Msg)
    FAIL("short arrays");
  for (int k = 0; k < 2; k++) {
    int32_t len;
    if (!rd(f, &len, 4)) FAIL("short tree string");
    char* txt = (char*)malloc((size_t)len + 1);
    if (len && !rd(f, txt, (size_t)len)) FAIL("short tree string");
    txt[len] = 0;
    int ni, nd;
    orc_tree* arr = parse_tree_array(txt, len, &ni, &nd);
    free(txt);
    if (k == 0) { if (!arr) FAIL("cannot parse var trees"); d->var_trees = arr; d->n_var_iter = ni; d->n_var_deg = nd; }
    else { d->chk_trees = arr; d->n_chk_iter = ni; d->n_chk_deg = nd; }
  }
  if (d->has_gen) {
    int32_t rc[2];
    if (!rd(f, rc, 8)) FAIL("short generator header");
    d->gen_rows = rc[0]; d->gen_cols = rc[1];
    d->gen = (uint8_t*)malloc((size_t)rc[0] * rc[1]);
    if (!rd(f, d->gen, (size_t)rc[0] * rc[1])) FAIL("short generator");
  }
  fclose(f); f = NULL;
  /* set_trees, src/LDPC_Code_LUT.cpp:120-160 */
  d->var_tree_idx_iter = (int32_t*)malloc(sizeof(int32_t) * (size_t)d->max_iters);
  int acc = 0;
  for (int i = 0; i < d->max_iters; i++) { acc += (d->reuse[i] == 0); d->var_tree_idx_iter[i] = acc - 1; }
  d->var_tree_idx_degree = (int32_t*)malloc(sizeof(int32_t) * (size_t)d->nvar);
  for (int v = 0; v < d->nvar; v++) {
    int idx = -1;
    for (int dd = 0; dd < d->n_var_deg && idx < 0; dd++)
      if (d->var_trees[dd].num_leaves == d->dv[v]) idx = dd;
    if (idx < 0) FAIL("no var tree for a variable-node degree");
    d->var_tree_idx_degree[v] = idx;
  }
  if (d->chk_trees && d->n_chk_iter > 0) {
    d->chk_tree_idx_degree = (int32_t*)malloc(sizeof(int32_t) * (size_t)d->nchk);
    for (int c = 0; c < d->nchk; c++) {
      int idx = -1;
      for (int dd = 0; dd < d->n_chk_deg && idx < 0; dd++)
        if (d->chk_trees[dd].num_leaves + 1 == d->dc[c]) idx = dd;
      if (idx < 0) FAIL("no chk tree for a check-node degree");
      d->chk_tree_idx_degree[c] = idx;
    }
  }
  /* chk_equ_idx (src/LDPC_Code_LUT.cpp:531-535): cn_msg_idx lists each check's edges in ascending VN
   * order (:513-527), so the VN owning each listed edge is the sorted row of H. */
  int32_t* edge_vn = (int32_t*)malloc(sizeof(int32_t) * (size_t)d->num_edges);
  int e = 0;
  for (int v = 0; v < d->nvar; v++) for (int k = 0; k < d->dv[v]; k++) edge_vn[e++] = v;
  d->chk_equ_idx = (int32_t*)malloc(sizeof(int32_t) * (size_t)d->num_edges);
  for (int i = 0; i < d->num_edges; i++) d->chk_equ_idx[i] = edge_vn[d->cn_msg_idx[i]];
  free(edge_vn);
  d->msgs = (int32_t*)malloc(sizeof(int32_t) * (size_t)d->num_edges);
  return d;
}

static void free_trees(orc_tree* a, int n) {
  if (!a) return;
  for (int i = 0; i < n; i++) {
    for (int k = 0; k < a[i].n_nodes; k++) { free(a[i].nodes[k].Q); free(a[i].nodes[k].child); }
    free(a[i].nodes);
  }
  free(a);
}
void orc_free(orc_decoder* d) {
  if (!d) return;
  free(d->dv); free(d->dc); free(d->cn_msg_idx); free(d->Nq_Msg); free(d->reuse); free(d->cha2msg);
  free(d->chk_equ_idx); free(d->qb_Cha); free(d->qb_Msg);
  free_trees(d->var_trees, d->n_var_iter * d->n_var_deg);
  free_trees(d->chk_trees, d->n_chk_iter * d->n_chk_deg);
  free(d->var_tree_idx_iter); free(d->var_tree_idx_degree); free(d->chk_tree_idx_degree);
  free(d->gen); free(d->msgs); free(d);
}
int orc_info(const orc_decoder* d, int what) {
  switch (what) {
    case ORC_NVAR: return d->nvar;
    case ORC_NCHK: return d->nchk;
    case ORC_NCHK_LIN_INDEP: return d->nchk_lin_indep;
    case ORC_NUM_EDGES: return d->num_edges;
    case ORC_MAX_ITERS: return d->max_iters;
    case ORC_NQ_CHA: return d->Nq_Cha;
    case ORC_MIN_LUT: return d->minLUT;
    case ORC_NINFO: return d->nvar - d->nchk_lin_indep; /* src/LDPC_Code_LUT.hpp:264 */
    case ORC_HAS_GEN: return d->has_gen;
    case ORC_N_ITER_SLOTS: return d->n_var_iter;
    case ORC_N_DEG_SLOTS: return d->n_var_deg;
  }
  return -1;
}
void orc_get_dv(const orc_decoder* d, int32_t* dst) { memcpy(dst, d->dv, sizeof(int32_t) * (size_t)d->nvar); }
void orc_get_dc(const orc_decoder* d, int32_t* dst) { memcpy(dst, d->dc, sizeof(int32_t) * (size_t)d->nchk); }
void orc_get_cn_msg_idx(const orc_decoder* d, int32_t* dst) {
  memcpy(dst, d->cn_msg_idx, sizeof(int32_t) * (size_t)d->num_edges);
}
void orc_get_nq_msg(const orc_decoder* d, int32_t* dst) {
  memcpy(dst, d->Nq_Msg, sizeof(int32_t) * (size_t)d->max_iters);
}
int orc_get_qb(const orc_decoder* d, int which, double* dst) {
  int n = which == 0 ? d->Nq_Cha - 1 : d->n_qb_msg;
  memcpy(dst, which == 0 ? d->qb_Cha : d->qb_Msg, sizeof(double) * (size_t)n);
  return n;
}

/* src/LDPC_Code_LUT.cpp:176-185 */
void orc_set_exit_conditions(orc_decoder* d, int max_iters, int psc, int pisc) {
  d->max_iters = max_iters; d->psc = psc; d->pisc = pisc;
}

/* src/common.cpp:120-128: count leading thresholds with x > b[i], stop at the first failure */
int orc_quant_nonlin(double x, const double* b, int nb) {
  int idx = 0;
  for (int i = 0; i < nb; i++) { if (x > b[i]) idx++; else break; }
  return idx;
}

/* src/LDPC_Code_LUT.cpp:207-221 */
void orc_quantise(const orc_decoder* d, const double* llr, int init_mode, uint8_t* cha, uint8_t* msg) {
  for (int v = 0; v < d->nvar; v++) {
    int c = orc_quant_nonlin(llr[v], d->qb_Cha, d->Nq_Cha - 1);
    cha[v] = (uint8_t)c;
    msg[v] = (uint8_t)(init_mode == ORC_INIT_CONT ? orc_quant_nonlin(llr[v], d->qb_Msg, d->n_qb_msg)
                                                  : d->cha2msg[c]);
  }
}

/* LUT_Tree_Node::var_msg_update, src/LUT_Tree.cpp:402-418.  `q` is the input queue, *pos its front:
 * a leaf of either type consumes the next input; an inner node forms a mixed-radix label from its
 * children (first child least significant) and looks it up in the half table with symmetry. */
static int node_var_update(const orc_tree* t, int n, const int* q, int* pos) {
  const orc_node* nd = &t->nodes[n];
  if (nd->type == NODE_MSG || nd->type == NODE_CHA) return q[(*pos)++];
  int label = 0, base = 1;
  for (int i = 0; i < nd->nchild; i++) {
    label += base * node_var_update(t, nd->child[i], q, pos);
    base *= t->nodes[nd->child[i]].K;
  }
  if (label < nd->qlen) return nd->Q[label];
  return nd->K - 1 - nd->Q[2 * nd->qlen - 1 - label];
}

/* LUT_Tree_Node::chk_msg_update, src/LUT_Tree.cpp:420-445 */
static int node_chk_update(const orc_tree* t, int n, const int* q, int* pos) {
  const orc_node* nd = &t->nodes[n];
  if (nd->type == NODE_MSG) return q[(*pos)++];
  int label = 0, base = 1, parity = 0;
  for (int i = 0; i < nd->nchild; i++) {
    int ls = node_chk_update(t, nd->child[i], q, pos);
    int res = t->nodes[nd->child[i]].K;
    if (ls < res / 2) { parity ^= 1; label += base * (res / 2 - 1 - ls); }
    else label += base * (ls - res / 2);
    base *= res / 2;
  }
  return parity == 1 ? nd->Q[label] : nd->K - 1 - nd->Q[label];
}

#define ORC_MAX_DEG 512

/* LUT_Tree::var_msg_update, src/LUT_Tree.cpp:774-790: output ii = tree over (inputs without ii) ++ llr */
static void tree_var_msg_update(const orc_tree* t, const int* in, int dv, int llr, int* out) {
  int q[ORC_MAX_DEG + 1];
  for (int ii = 0; ii < dv; ii++) {
    int n = 0;
    for (int k = 0; k < dv; k++) if (k != ii) q[n++] = in[k];
    q[n++] = llr;
    int pos = 0;
    out[ii] = node_var_update(t, 0, q, &pos);
  }
}
/* LUT_Tree::chk_msg_update, src/LUT_Tree.cpp:792-807 */
static void tree_chk_msg_update(const orc_tree* t, const int* in, int dc, int* out) {
  int q[ORC_MAX_DEG + 1];
  for (int ii = 0; ii < dc; ii++) {
    int n = 0;
    for (int k = 0; k < dc; k++) if (k != ii) q[n++] = in[k];
    int pos = 0;
    out[ii] = node_chk_update(t, 0, q, &pos);
  }
}
/* LUT_Tree::dec_update, src/LUT_Tree.cpp:809-820 */
static int tree_dec_update(const orc_tree* t, const int* in, int dv, int llr) {
  int q[ORC_MAX_DEG + 1];
  for (int k = 0; k < dv; k++) q[k] = in[k];
  q[dv] = llr;
  int pos = 0;
  return node_var_update(t, 0, q, &pos);
}

/* syndrome_check(const bvec&), src/LDPC_Code_LUT.cpp:455-469 */
int orc_syndrome_ok(const orc_decoder* d, const uint8_t* b) {
  int e = 0;
  for (int cc = 0; cc < d->nchk; cc++) {
    int synd = 0;
    for (int i = 0; i < d->dc[cc]; i++) if (b[d->chk_equ_idx[e + i]]) synd++;
    e += d->dc[cc];
    if (synd & 1) return 0;
  }
  return 1;
}

/* syndrome_check(int Nq_Msg, bvec&), src/LDPC_Code_LUT.cpp:437-452: every VN's outgoing messages
 * must agree in sign, then the resulting word must satisfy H. */
static int unanimous_syndrome_ok(const orc_decoder* d, int Nq, uint8_t* b) {
  int e = 0, nz = Nq / 2;
  for (int v = 0; v < d->nvar; v++) {
    int bit = d->msgs[e] < nz;
    for (int k = 1; k < d->dv[v]; k++) if ((d->msgs[e + k] < nz) != bit) return 0;
    e += d->dv[v];
    b[v] = (uint8_t)bit;
  }
  return orc_syndrome_ok(d, b);
}

/* chk_update_minsum, src/LDPC_Code_LUT.cpp:355-402 */
static void chk_update_minsum(orc_decoder* d, int cc, int edge_idx, int iter) {
  int nz = d->Nq_Msg[iter] / 2, min1 = nz, min2 = nz, min_idx = 0, sign_prod = 0, tmp;
  int dc = d->dc[cc];
  const int32_t* idx = d->cn_msg_idx + edge_idx;
  for (int k = 0; k < dc; k++) {
    int m = d->msgs[idx[k]];
    if (m < nz) { sign_prod ^= 1; tmp = nz - 1 - m; } else tmp = m - nz;
    if (tmp < min1) { min2 = min1; min1 = tmp; min_idx = k; }
    else if (tmp < min2) min2 = tmp;
  }
  for (int k = 0; k < dc; k++) {
    tmp = (k == min_idx) ? min2 : min1;
    int sign_msg = (d->msgs[idx[k]] < nz) ? (sign_prod ^ 1) : sign_prod;
    d->msgs[idx[k]] = sign_msg ? nz - 1 - tmp : nz + tmp;
  }
}

static void trace_put(const orc_decoder* d, int kind, int iter, uint8_t* tm, int32_t* tk, int32_t* ti,
                      int max_rec, int32_t* n_rec) {
  if (!tm || !n_rec || *n_rec >= max_rec) return;
  uint8_t* dst = tm + (size_t)(*n_rec) * d->num_edges;
  for (int e = 0; e < d->num_edges; e++) dst[e] = (uint8_t)d->msgs[e];
  if (tk) tk[*n_rec] = kind;
  if (ti) ti[*n_rec] = iter;
  (*n_rec)++;
}

/* lut_decode, src/LDPC_Code_LUT.cpp:259-353 */
int orc_lut_decode(orc_decoder* d, const uint8_t* cha, const uint8_t* msg, uint8_t* out,
                   uint8_t* tm, int32_t* tk, int32_t* ti, int max_rec, int32_t* n_rec) {
  int in[ORC_MAX_DEG], res[ORC_MAX_DEG];
  if (n_rec) *n_rec = 0;
  for (int v = 0; v < d->nvar; v++) out[v] = cha[v] < d->Nq_Cha / 2;            /* :275 */
  if (d->pisc && orc_syndrome_ok(d, out)) return 0;                             /* :277 */
  int e = 0;
  for (int v = 0; v < d->nvar; v++) for (int k = 0; k < d->dv[v]; k++) d->msgs[e++] = msg[v]; /* :284 */
  trace_put(d, ORC_TRACE_INIT, -1, tm, tk, ti, max_rec, n_rec);
  for (int ii = 0; ii < d->max_iters; ii++) {
    e = 0;
    for (int cc = 0; cc < d->nchk; cc++) {                                      /* :304-310 */
      if (d->minLUT) chk_update_minsum(d, cc, e, ii);
      else { /* chk_update_lut, :416-426 */
        int dc = d->dc[cc];
        for (int k = 0; k < dc; k++) in[k] = d->msgs[d->cn_msg_idx[e + k]];
        const orc_tree* t = &d->chk_trees[d->var_tree_idx_iter[ii] * d->n_chk_deg + d->chk_tree_idx_degree[cc]];
        tree_chk_msg_update(t, in, dc, res);
        for (int k = 0; k < dc; k++) d->msgs[d->cn_msg_idx[e + k]] = res[k];
      }
      e += d->dc[cc];
    }
    trace_put(d, ORC_TRACE_CN, ii, tm, tk, ti, max_rec, n_rec);
    if (ii != d->max_iters - 1) {                                               /* :320-330 */
      e = 0;
      for (int v = 0; v < d->nvar; v++) { /* var_update_lut, :404-414 */
        int dv = d->dv[v];
        for (int k = 0; k < dv; k++) in[k] = d->msgs[e + k];
        const orc_tree* t = &d->var_trees[d->var_tree_idx_iter[ii] * d->n_var_deg + d->var_tree_idx_degree[v]];
        tree_var_msg_update(t, in, dv, cha[v], res);
        for (int k = 0; k < dv; k++) d->msgs[e + k] = res[k];
        e += dv;
      }
      if (d->psc && unanimous_syndrome_ok(d, d->Nq_Msg[ii + 1], out)) return ii + 1;
    }
    trace_put(d, ORC_TRACE_VN, ii, tm, tk, ti, max_rec, n_rec);                 /* :331-337 */
  }
  e = 0;
  for (int v = 0; v < d->nvar; v++) { /* decision, :340-344; dec_update_lut :428-434 */
    int dv = d->dv[v];
    for (int k = 0; k < dv; k++) in[k] = d->msgs[e + k];
    const orc_tree* t = &d->var_trees[d->var_tree_idx_iter[d->max_iters - 1] * d->n_var_deg + d->var_tree_idx_degree[v]];
    out[v] = tree_dec_update(t, in, dv, cha[v]) < 1;
    e += dv;
  }
  return orc_syndrome_ok(d, out) ? d->max_iters : -d->max_iters;                /* :346-349 */
}

void orc_lut_decode_batch(orc_decoder* d, int n_frames, const uint8_t* cha, const uint8_t* msg,
                          uint8_t* out, int32_t* ret) {
  for (int f = 0; f < n_frames; f++) {
    size_t o = (size_t)f * d->nvar;
    ret[f] = orc_lut_decode(d, cha + o, msg + o, out + o, NULL, NULL, NULL, 0, NULL);
  }
}

/* LDPC_Generator_Systematic::encode, itpp/itpp/comm/ldpc.cpp:1038: output = [input, G*input] */
int orc_encode(const orc_decoder* d, const uint8_t* info, uint8_t* cw) {
  if (!d->has_gen) return -1;
  int K = d->gen_cols;
  for (int k = 0; k < K; k++) cw[k] = info[k];
  for (int r = 0; r < d->gen_rows; r++) {
    int acc = 0;
    const uint8_t* row = d->gen + (size_t)r * K;
    for (int k = 0; k < K; k++) acc ^= row[k] & info[k];
    cw[K + r] = (uint8_t)acc;
  }
  return 0;
}
