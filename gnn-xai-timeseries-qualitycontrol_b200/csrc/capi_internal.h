// Shared between the translation units that implement extern "C" entry points.
#pragma once
#include <string>

#include "device_model.cuh"

struct gnnqc_model {
  gnnqc::DeviceModel dm;
};

// Records `msg` as the calling thread's gnnqc_last_error() and returns `code`.
int gnnqc_internal_fail(int code, const std::string& msg);
