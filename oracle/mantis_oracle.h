/*
 * mantis_oracle.h -- CPU oracle for the mantis localization hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a plain-C restatement of the arithmetic the
 * reference (k-sheridan/mantis, mantis3 generation) performs per frame, including
 * the OpenCV 4.13 fixed-point kernels it calls.  It exists to check the CUDA
 * library (libmantis_cuda.so) and to be timed as the CPU baseline.  Nothing in the
 * product path may include, link or call it.
 *
 * Parity pin: every function here is checked against cv2 4.13.0 (the OpenCV the
 * reference's calls resolve to in this environment) on test/grid1.jpg,
 * test/lena.jpg and seeded noise by tests/test_oracle_*.py, with golden digests
 * committed under tests/golden/.  The reference itself has no assertions or
 * recorded outputs for these stages.
 *
 * Each function cites the reference call site it restates (file:line relative to
 * the reference tree).
 */
#ifndef MANTIS_ORACLE_H
#define MANTIS_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- image stage: include/mantis3/QuadDetection.h:209-214 ---- */
void mo_bgr2gray(const uint8_t* bgr, int w, int h, int stride, uint8_t* gray);
void mo_blur3(const uint8_t* src, int w, int h, uint8_t* dst);
void mo_sobel3(const uint8_t* src, int w, int h, int16_t* dx, int16_t* dy);
void mo_canny(const uint8_t* src, int w, int h, int low, int high, uint8_t* dst);
void mo_dilate3(const uint8_t* src, int w, int h, int iters, uint8_t* dst);
void mo_erode3(const uint8_t* src, int w, int h, int iters, uint8_t* dst);
/* whole chain of detectQuadrilaterals up to the contour input */
void mo_edge_stage(const uint8_t* bgr, int w, int h, int stride, int canny_low,
                   uint8_t* canny_out /* may be NULL */, uint8_t* morph_out);

#ifdef __cplusplus
}
#endif
#endif
