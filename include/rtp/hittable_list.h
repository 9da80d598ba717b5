// hittable_list.h -- same name as the reference header (src/hittable_list.h); the API lives in scene_api.h
#include "scene_api.h"
