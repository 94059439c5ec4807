// frontend_api.cpp — C-ABI face of the scene-file front-end (include/akari_b200.h, "scene files").
#include <exception>
#include <string>

#include "scene_file.h"

namespace akb {
akb_status set_error(akb_status st, const std::string &msg); // api.cu
}

struct akb_scene_file {
    std::unique_ptr<akb_host::SceneFile> sf;
};

namespace {
template <class F> akb_status guarded(const char *who, F &&body) {
    try {
        body();
        return AKB_OK;
    } catch (const std::bad_alloc &) {
        return akb::set_error(AKB_OOM, std::string(who) + ": out of host memory");
    } catch (const std::exception &e) {
        return akb::set_error(AKB_INVALID_ARG, std::string(who) + ": " + e.what());
    }
}
} // namespace

extern "C" {

akb_status akb_scene_file_load(const char *path, akb_scene_file **out_file) {
    if (!path || !out_file) return akb::set_error(AKB_INVALID_ARG, "akb_scene_file_load: null argument");
    *out_file = nullptr;
    return guarded("akb_scene_file_load", [&] {
        auto h = std::make_unique<akb_scene_file>();
        h->sf = akb_host::load_scene_file(path);
        *out_file = h.release();
    });
}

const akb_scene_desc *akb_scene_file_desc(const akb_scene_file *file) { return file ? &file->sf->desc : nullptr; }

akb_status akb_scene_file_params(const akb_scene_file *file, akb_render_params *out_params) {
    if (!file || !out_params) return akb::set_error(AKB_INVALID_ARG, "akb_scene_file_params: null argument");
    *out_params = file->sf->params;
    return AKB_OK;
}

const char *akb_scene_file_output(const akb_scene_file *file) { return file ? file->sf->output.c_str() : ""; }

akb_status akb_scene_file_destroy(akb_scene_file *file) {
    delete file;
    return AKB_OK;
}

akb_status akb_write_png(const char *path, const uint8_t *rgb8, int32_t width, int32_t height) {
    if (!path || !rgb8 || width <= 0 || height <= 0) return akb::set_error(AKB_INVALID_ARG, "akb_write_png: bad argument");
    return guarded("akb_write_png", [&] { akb_host::write_png_rgb8(path, rgb8, width, height); });
}

akb_status akb_write_pfm(const char *path, const float *rgbw, int32_t width, int32_t height) {
    if (!path || !rgbw || width <= 0 || height <= 0) return akb::set_error(AKB_INVALID_ARG, "akb_write_pfm: bad argument");
    return guarded("akb_write_pfm", [&] { akb_host::write_pfm_rgb(path, rgbw, width, height); });
}

} // extern "C"
