// Pairwise Rosenbrock, float32: compiled kernel shapes.  Part 2 of 2: the shape list is split over
// several translation units only so that they compile in parallel.
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(RosenbrockModel, HMCB_MODEL_ROSENBROCK, float, HMCB_F32, UNITM, EPL, LPC)
S(1, 4, 1) S(1, 20, 1) S(1, 50, 1) S(1, 28, 4) S(1, 4, 32) S(0, 10, 1)
