// Pairwise Rosenbrock, float64: compiled kernel shapes.  Part 3 of 4: the shape list is split over
// several translation units only so that they compile in parallel.
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(RosenbrockModel, HMCB_MODEL_ROSENBROCK, double, HMCB_F64, UNITM, EPL, LPC)
S(1, 10, 1) S(1, 32, 1) S(1, 26, 2) S(1, 8, 32) S(0, 20, 1) S(0, 4, 32)
