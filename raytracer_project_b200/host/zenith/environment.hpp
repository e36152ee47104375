// zenith/environment.hpp — EnvironmentSettings of the host scene API (environment.hpp:8-77).
// Same fields and defaults as the reference; consumed by the CUDA miss shader that restates
// camera::get_background_color (camera.hpp:828-925).
#pragma once
#include <string>

#include "materials.hpp"

namespace zenith {

struct EnvironmentSettings {
    enum Mode { PHYSICAL_SUN, HDR_MAP, SOLID_COLOR };
    Mode _mode = PHYSICAL_SUN;
    bool needs_ui_sync = false;
    Mode mode() const { return _mode; }
    void set_mode(Mode m) {
        if (_mode != m) { _mode = m; needs_ui_sync = true; }
    }
    color background_color = color(0.0, 0.0, 0.0);
    std::string current_hdr_name = "None";
    std::string current_hdr_path = "assets/hdr_maps/";
    double intensity = 1.0;
    double hdri_rotation = 0.0, hdri_tilt = 0.0, hdri_roll = 0.0;  // radians
    shared_ptr<image_texture> hdr_texture = nullptr;
    // environment.hpp:46-69: empty path => SOLID_COLOR black; otherwise HDR_MAP
    void load_hdr(const std::string& path) {
        if (path.empty()) {
            set_mode(SOLID_COLOR);
            background_color = color(0, 0, 0);
            current_hdr_name = "None (Black)";
            return;
        }
        hdr_texture = make_shared<image_texture>(path.c_str(), true);
        size_t slash = path.find_last_of("/\\");
        current_hdr_name = slash == std::string::npos ? path : path.substr(slash + 1);
        current_hdr_path = path;
        set_mode(HDR_MAP);
    }
    vec3 sun_direction = unit_vector(vec3(1.0, 0.5, -0.5));
    color sun_color = color(1.0, 1.0, 1.0);
    bool auto_sun_color = true;
    double sun_intensity = 1.0;
    double sun_size = 1.0;
};

}  // namespace zenith
