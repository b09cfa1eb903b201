// yaml-cpp is not installed in this image: the reference's `#include "yaml-cpp/yaml.h"` lands here and gets the repo's
// own YAML reader, which mirrors the yaml-cpp calls the reference makes (oracle/Makefile, target _ref/hypre_app_ref).
#pragma once
#include "yaml_lite.h"
