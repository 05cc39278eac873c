// Dense Gaussian, float64: compiled kernel shapes.
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(GaussianModel, HMCB_MODEL_GAUSSIAN, double, HMCB_F64, UNITM, EPL, LPC)
S(1, 2, 1) S(1, 4, 1) S(1, 10, 1) S(1, 20, 1) S(1, 26, 2) S(1, 14, 4) S(1, 28, 4) S(1, 8, 8) S(1, 16, 8) S(1, 4, 32) S(1, 8, 32)
S(0, 4, 1) S(0, 10, 1) S(0, 14, 4) S(0, 28, 4) S(0, 8, 32)
