/* TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path.
 *
 * Plain-C restatement of the BP comparison decoder ber_sim runs for a [BP] section
 * (src/LDPC_BER_Sim.cpp:157-244): IT++'s LDPC_Code::bp_decode over integer LLRs (QLLR) with the
 * LLR_calc_unit of the reference's vendored, modified IT++ (saturating qsum, total resolution Dint4).
 *
 * Parity status: PINNED against outputs of the reference itself (oracle/_ref/lutldpc_ref bpfixture:
 * channel LLRs, to_qllr of them, bp_decode's soft output and return value; tests/golden/bp_*.fix.xz).
 *
 * The graph is given the way the LUT decoder's bundle holds it (src/LDPC_Code_LUT.cpp:513-528): edges
 * numbered VN-major (variable i owns edges off[i] .. off[i]+dv[i]-1, ascending check index -- the order of
 * C(j0 + cmax*i), itpp/itpp/comm/ldpc.cpp:1696-1701) and cn_msg_idx listing, check by check, the edge ids in
 * ascending variable order (the order of V(j + ncheck*i0) / jind, :1705-1726).  Edge (i, k) is the reference's
 * mvc[i + k*nvar]; position ip of check j is its mcv[j + ip*ncheck].
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef int32_t QLLR;
#define QLLR_MAX_CONST (INT32_MAX >> 4) /* itpp/itpp/comm/llr.h:61 */

typedef struct {
  int d1, d2, d3, d4;
  QLLR qllr_max; /* init_llr_bounds, itpp/itpp/comm/llr.cpp:74-80 */
  QLLR* table;   /* construct_logexp_table, :82-92 */
} llrcalc;

/* LLR_calc_unit::to_qllr, itpp/itpp/comm/llr.h:266-279 (positive overflow returns QLLR_MAX, not qllr_max) */
static QLLR to_qllr(const llrcalc* c, double l) {
  double qmax_d = (double)c->qllr_max / (double)(1 << c->d1);
  if (l > qmax_d) return QLLR_MAX_CONST;
  if (l < -qmax_d) return -c->qllr_max;
  return (QLLR)floor(0.5 + (double)(1 << c->d1) * l);
}
static void llrcalc_init(llrcalc* c, int d1, int d2, int d3, int d4) {
  c->d1 = d1; c->d2 = d2; c->d3 = d3; c->d4 = d4;
  c->qllr_max = (1 << (d4 - 1)) - 1;
  c->table = (QLLR*)malloc(sizeof(QLLR) * (d2 > 0 ? d2 : 1));
  for (int i = 0; i < d2; i++) {
    double x = pow(2.0, (double)(d3 - d1)) * i;
    c->table[i] = to_qllr(c, log(1 + exp(-x)));
  }
}
/* llr.h:282-298 */
static QLLR logexp(const llrcalc* c, QLLR x) {
  int ind = x >> c->d3;
  if (ind >= c->d2) return 0;
  return c->table[ind];
}
/* llr.h:320-329 */
static QLLR qsum(const llrcalc* c, QLLR a, QLLR b) {
  QLLR s = a + b;
  if (s > c->qllr_max) return c->qllr_max;
  if (s < -c->qllr_max) return -c->qllr_max;
  return s;
}
/* LLR_calc_unit::Boxplus, llr.cpp:148-185 */
static QLLR boxplus(const llrcalc* c, QLLR a, QLLR b) {
  QLLR a_abs = a > 0 ? a : -a, b_abs = b > 0 ? b : -b;
  QLLR minabs = a_abs > b_abs ? b_abs : a_abs;
  QLLR term1 = a > 0 ? (b > 0 ? minabs : -minabs) : (b > 0 ? -minabs : minabs);
  if (c->d2 == 0) {
    if (term1 > c->qllr_max) return c->qllr_max;
    if (term1 < -c->qllr_max) return -c->qllr_max;
    return term1;
  }
  QLLR apb = a + b, amb = a - b;
  QLLR r = term1 + logexp(c, apb > 0 ? apb : -apb) - logexp(c, amb > 0 ? amb : -amb);
  if (r > c->qllr_max) return c->qllr_max;
  if (r < -c->qllr_max) return -c->qllr_max;
  return r;
}

void bp_oracle_to_qllr(int d1, int d4, const double* llr, int n, int32_t* out) {
  llrcalc c;
  c.d1 = d1; c.d4 = d4; c.qllr_max = (1 << (d4 - 1)) - 1;
  for (int i = 0; i < n; i++) out[i] = to_qllr(&c, llr[i]);
}
void bp_oracle_logexp_table(int d1, int d2, int d3, int d4, int32_t* out) {
  llrcalc c;
  llrcalc_init(&c, d1, d2, d3, d4);
  memcpy(out, c.table, sizeof(QLLR) * d2);
  free(c.table);
}

/* syndrome_check(const QLLRvec&), itpp/itpp/comm/ldpc.cpp:1631-1652 */
static int syndrome_ok(int nchk, const int32_t* dc, const int32_t* cn_off, const int32_t* cn_var, const QLLR* llr) {
  for (int j = 0; j < nchk; j++) {
    int synd = 0;
    for (int i = 0; i < dc[j]; i++) synd += llr[cn_var[cn_off[j] + i]] < 0;
    if (synd & 1) return 0;
  }
  return 1;
}

/* LDPC_Code::bp_decode, itpp/itpp/comm/ldpc.cpp:1398-1622.  Returns +iter (valid) / -iter, 0 if pisc passes;
 * -9999 on a graph the reference refuses (check degree < 2 or > 200, variable degree 0). */
int bp_oracle_decode(int nvar, int nchk, const int32_t* dv, const int32_t* dc, const int32_t* cn_msg_idx, int max_iters,
                     int psc, int pisc, int d1, int d2, int d3, int d4, const int32_t* LLRin, int32_t* LLRout) {
  llrcalc lc;
  llrcalc_init(&lc, d1, d2, d3, d4);
  int E = 0, maxdc = 0;
  int32_t* off = (int32_t*)malloc(sizeof(int32_t) * (nvar + 1));
  for (int i = 0; i < nvar; i++) { off[i] = E; E += dv[i]; }
  off[nvar] = E;
  int32_t* cn_off = (int32_t*)malloc(sizeof(int32_t) * (nchk + 1));
  int acc = 0;
  for (int j = 0; j < nchk; j++) { cn_off[j] = acc; acc += dc[j]; if (dc[j] > maxdc) maxdc = dc[j]; }
  cn_off[nchk] = acc;
  int32_t* edge_var = (int32_t*)malloc(sizeof(int32_t) * (E > 0 ? E : 1));
  for (int i = 0; i < nvar; i++) for (int k = 0; k < dv[i]; k++) edge_var[off[i] + k] = i;
  int32_t* cn_var = (int32_t*)malloc(sizeof(int32_t) * (E > 0 ? E : 1));
  for (int e = 0; e < E; e++) cn_var[e] = edge_var[cn_msg_idx[e]];
  QLLR* msg = (QLLR*)malloc(sizeof(QLLR) * (E > 0 ? E : 1)); /* mvc and mcv share storage: each edge has one of each */
  QLLR* m = (QLLR*)malloc(sizeof(QLLR) * (maxdc + 1));
  QLLR* ml = (QLLR*)malloc(sizeof(QLLR) * (maxdc + 1));
  QLLR* mr = (QLLR*)malloc(sizeof(QLLR) * (maxdc + 1));
  int ret = -9999, bad = 0;
  for (int j = 0; j < nchk; j++) if (dc[j] < 2 || dc[j] > 200) bad = 1; /* :1439-1443, :1523-1527 */
  for (int i = 0; i < nvar; i++) if (dv[i] < 1) bad = 1;                  /* :1563-1564 */
  if (bad) goto done;

  if (pisc && syndrome_ok(nchk, dc, cn_off, cn_var, LLRin)) { /* :1408-1411 */
    memcpy(LLRout, LLRin, sizeof(QLLR) * nvar);
    ret = 0;
    goto done;
  }
  for (int i = 0; i < nvar; i++) for (int k = 0; k < dv[i]; k++) msg[off[i] + k] = LLRin[i]; /* :1422-1428 */

  int valid = 0, iter = 0;
  do {
    iter++;
    /* step 1, check to variable, :1436-1556 */
    for (int j = 0; j < nchk; j++) {
      const int32_t* idx = cn_msg_idx + cn_off[j];
      int n = dc[j];
      for (int k = 0; k < n; k++) m[k] = msg[idx[k]];
      if (n == 6) { /* the one hard-coded degree whose tree differs from the general recursion, :1497-1521 */
        QLLR m01 = boxplus(&lc, m[0], m[1]), m23 = boxplus(&lc, m[2], m[3]), m45 = boxplus(&lc, m[4], m[5]);
        QLLR m03 = boxplus(&lc, m01, m23), m25 = boxplus(&lc, m23, m45), m0145 = boxplus(&lc, m01, m45);
        msg[idx[0]] = boxplus(&lc, m[1], m25);
        msg[idx[1]] = boxplus(&lc, m[0], m25);
        msg[idx[2]] = boxplus(&lc, m0145, m[3]);
        msg[idx[3]] = boxplus(&lc, m0145, m[2]);
        msg[idx[4]] = boxplus(&lc, m03, m[5]);
        msg[idx[5]] = boxplus(&lc, m03, m[4]);
      } else { /* degrees 2..5 (:1444-1496) compute exactly this recursion (:1529-1553) */
        int nodes = n - 1;
        ml[0] = m[0];
        mr[0] = m[nodes];
        for (int i = 1; i < nodes; i++) {
          ml[i] = boxplus(&lc, ml[i - 1], m[i]);
          mr[i] = boxplus(&lc, mr[i - 1], m[nodes - i]);
        }
        msg[idx[0]] = mr[nodes - 1];
        msg[idx[nodes]] = ml[nodes - 1];
        for (int i = 1; i < nodes; i++) msg[idx[i]] = boxplus(&lc, ml[i - 1], mr[nodes - 1 - i]);
      }
    }
    /* step 2, variable to check, :1558-1611 */
    for (int i = 0; i < nvar; i++) {
      QLLR* e = msg + off[i];
      switch (dv[i]) {
        case 1: {
          QLLR m0 = e[0];
          e[0] = LLRin[i];
          LLRout[i] = qsum(&lc, LLRin[i], m0);
          break;
        }
        case 2: {
          QLLR m0 = e[0], m1 = e[1];
          e[0] = qsum(&lc, LLRin[i], m1);
          e[1] = qsum(&lc, LLRin[i], m0);
          LLRout[i] = qsum(&lc, e[1], m1);
          break;
        }
        default: { /* case 3 (:1583-1595) is the same arithmetic as the differential update (:1596-1609) */
          QLLR t = LLRin[i];
          for (int k = 0; k < dv[i]; k++) t = qsum(&lc, t, e[k]);
          LLRout[i] = t;
          for (int k = 0; k < dv[i]; k++) e[k] = qsum(&lc, t, -e[k]);
        }
      }
    }
    if (psc && syndrome_ok(nchk, dc, cn_off, cn_var, LLRout)) { valid = 1; break; } /* :1613-1616 */
  } while (iter < max_iters);
  ret = valid ? iter : -iter; /* :1621 */
done:
  free(off); free(cn_off); free(edge_var); free(cn_var); free(msg); free(m); free(ml); free(mr); free(lc.table);
  return ret;
}
