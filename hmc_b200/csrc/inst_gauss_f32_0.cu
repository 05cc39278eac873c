// Dense Gaussian, float32: compiled kernel shapes.  Part 1 of 2: the shape list is split over
// several translation units only so that they compile in parallel.
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(GaussianModel, HMCB_MODEL_GAUSSIAN, float, HMCB_F32, UNITM, EPL, LPC)
S(1, 4, 1) S(1, 10, 1) S(1, 28, 4) S(1, 4, 32) S(1, 8, 32) S(0, 14, 4)
