// Dense Gaussian, float64: compiled kernel shapes.  Part 4 of 4: the shape list is split over
// several translation units only so that they compile in parallel.
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(GaussianModel, HMCB_MODEL_GAUSSIAN, double, HMCB_F64, UNITM, EPL, LPC)
S(1, 2, 1) S(1, 20, 1) S(1, 16, 8) S(0, 10, 1) S(0, 8, 32)
