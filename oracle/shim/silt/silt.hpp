#pragma once
#include <silt/core/types.hpp>
#include <silt/core/shape.hpp>
#include <silt/core/tensor.hpp>
#include <silt/core/error.hpp>
#include <silt/op/common.hpp>
