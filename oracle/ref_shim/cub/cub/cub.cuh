/* Build shim (oracle/Makefile.refcuda): the reference includes its vendored CUB 1.8.0 as
 * <cub/cub/cub.cuh>, which clashes with the CUB that ships inside the CUDA 12.9 toolkit.
 * This directory sits in front of the reference's include path and hands the toolkit's CUB over. */
#pragma once
#include <cub/cub.cuh>
