/* TEST INFRASTRUCTURE ONLY — the CPU oracle for the polar encode / SC / SCL decode path.
 *
 * Plain-C restatement of the algorithms in dim49v/polarCodes (reference files cited per
 * function in polar_oracle.c).  Parity status: PINNED — every function here is checked
 * bit-for-bit against the reference's own code compiled in place (oracle/_ref, built by
 * oracle/Makefile from /root/reference) in tests/test_oracle_vs_ref.py, and against the
 * committed golden vectors under tests/golden/ (generated from oracle/_ref by
 * oracle/gen_golden.py) in tests/test_oracle_golden.py.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (libpolar_b200.so) never links or calls it.
 */
#ifndef POLAR_ORACLE_H
#define POLAR_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

double po_sigma(int n, int k, double snr_db);
void po_inf_indexes_awgn(int n, int k, double sigma, int* out);
void po_inf_indexes_bec(int n, int k, int* out);
void po_frozen_mask(int n, const int* inf, int ninf, int* mask);
void po_add_frozen_bits(const int* xinf, int* x, int n, const int* inf, int ninf);
void po_encode(int n, const int* u, int* x, int with_shuffle);
unsigned po_crc8(const int* bits, int len);
unsigned po_crc16(const int* bits, int len);
void po_append_crc(int* xinf, int k, int crc);
void po_scms_decode(int n, const double* y, const int* mask, int* sol);
void po_sc_decode(int n, const double* y, const int* mask, double sigma, int* sol);
/* returns 0 on success, -1 on allocation failure / bad L */
int po_scl_decode(int n, int L, int k, int crc, const double* y, const int* mask, int* sol);
/* batch helpers: decoder 1 SC, 2 SCMinsum, 3 SCLMinsum; sol stride = n (1,2) or k+crc (3) */
int po_decode_batch(int decoder, int n, int L, int k, int crc, double sigma, const double* y,
                    const int* mask, int* sol, int frames);
/* 8-bit min-sum SC (the product's PCB_PREC_INT8 mode; see polar_oracle.c) */
int po_q8(double y, int qfraction);
int po_f8(int r1, int r2);
int po_g8(int r1, int r2, int b);
int po_sc8_decode_batch(int n, int qfraction, const double* y, const int* mask, int* sol, int frames);
double po_time_decode(int decoder, int n, int L, int k, int crc, double sigma, const double* y,
                      const int* mask, int frames, int reps);

#ifdef __cplusplus
}
#endif
#endif
