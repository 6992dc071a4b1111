// er_commands.hpp — the command layer of the reference host behind the unchanged wire protocol:
//   parse_input_command   src/main.cpp:13-187   (a command is a CLI string; the messages that carry its data follow)
//   CommandManager        src/CommandManager.cpp:8-236 (JSON -> Camera / Texture / Material / RenderParameters),
//                         :250-504 (handlers; every load and `--start` answers STATUS "ok")
// over er::RenderingManager (include/er_manager.hpp), i.e. the C ABI, instead of SYCL.
#pragma once
#include <iomanip>
#include <memory>
#include <sstream>

#include "er_hostscene.hpp"
#include "er_message.hpp"
#include "er_objloader.hpp"

namespace erh {

// str_to_argv (src/main.cpp:13-24): whitespace-separated words, double quotes group and backslash escapes
inline std::vector<std::string> split_command(const std::string& str) {
    std::vector<std::string> out;
    std::istringstream iss(str);
    std::string s;
    while (iss >> std::quoted(s)) out.push_back(s);
    return out;
}

struct InputCommand {  // what program_options leaves in `vm` (src/main.cpp:48-83)
    std::vector<std::string> flags;  // options present, without the leading dashes
    std::vector<std::string> path;   // --path tokens
    std::string get_pass, output;
    bool has(const char* f) const {
        for (const auto& s : flags)
            if (s == f) return true;
        return false;
    }
};

inline InputCommand parse_command(const std::string& command_str) {
    static const char* const kFlags[] = {"load_config", "load_texture", "load_object", "load_camera", "load_hdri", "load_brdf_material",
                                         "load_osl_material", "start", "pause", "abort", "help", "recompute_normals", "mirror_x",
                                         "mirror_y", "get_info", "get_sycl_info"};
    InputCommand ic;
    const std::vector<std::string> argv = split_command(command_str);
    for (size_t i = 0; i < argv.size(); i++) {
        const std::string& a = argv[i];
        if (a.compare(0, 2, "--") != 0) throw std::runtime_error("too many positional options: " + a);
        std::string name = a.substr(2), value;
        bool has_value = false;
        const size_t eq = name.find('=');
        if (eq != std::string::npos) {
            value = name.substr(eq + 1);
            name = name.substr(0, eq);
            has_value = true;
        }
        if (name == "get_pass" || name == "output") {
            if (!has_value) {
                if (i + 1 >= argv.size()) throw std::runtime_error("the required argument for option '--" + name + "' is missing");
                value = argv[++i];
            }
            (name == "get_pass" ? ic.get_pass : ic.output) = value;
            ic.flags.push_back(name);
        } else if (name == "path") {  // multitoken
            if (has_value) ic.path.push_back(value);
            while (i + 1 < argv.size() && argv[i + 1].compare(0, 2, "--") != 0) ic.path.push_back(argv[++i]);
            if (ic.path.empty()) throw std::runtime_error("the required argument for option '--path' is missing");
            ic.flags.push_back(name);
        } else {
            bool known = false;
            for (const char* f : kFlags) known = known || name == f;
            if (!known) throw std::runtime_error("unrecognised option '--" + name + "'");
            ic.flags.push_back(name);
        }
    }
    return ic;
}

// ---- JSON -> records (src/CommandManager.cpp:8-112) ------------------------------------------------
inline ErCamera parse_camerajson(const erj::Value& j, ErCamera c) {
    const erj::Value& p = j.at("position");
    const erj::Value& r = j.at("rotation");
    c.aperture = (float)j.at("aperture").as_number();
    c.bokeh = j.at("bokeh").as_bool() ? 1 : 0;
    c.focusDistance = (float)j.at("focus_distance").as_number();
    c.focalLength = (float)j.at("focal_length").as_number();
    c.sensorWidth = (float)j.at("sensor_width").as_number();
    c.sensorHeight = (float)j.at("sensor_height").as_number();
    c.position[0] = (float)p.at("x").as_number(), c.position[1] = (float)p.at("y").as_number(), c.position[2] = (float)p.at("z").as_number();
    c.rotation[0] = (float)r.at("x").as_number(), c.rotation[1] = (float)r.at("y").as_number(), c.rotation[2] = (float)r.at("z").as_number();
    return c;
}

inline Texture parse_texturejson(const erj::Value& meta, const float* data, size_t count) {
    const int w = (int)meta.at("width").as_int64(), h = (int)meta.at("height").as_int64(), ch = (int)meta.at("channels").as_int64();
    const std::string& cs = meta.at("color_space").as_string();
    Texture::CS c = Texture::CS::LINEAR;  // an unknown colour space logs a warning in the reference and is left undefined
    if (cs == "sRGB") c = Texture::CS::sRGB;
    else if (cs != "LINEAR") fprintf(stderr, "Texture metadata colorspace not recognized\n");
    return Texture(meta.at("name").as_string(), w, h, ch, data, count, Texture::Filter::NO_FILTER, c);  // TCP textures are never filtered (:41)
}

inline Material parse_materialjson(const erj::Value& j) {
    Material mtl;
    auto rgb = [](const erj::Value& v, float* out) {
        out[0] = (float)v.at("r").as_number(), out[1] = (float)v.at("g").as_number(), out[2] = (float)v.at("b").as_number();
    };
    if (j.contains("name")) mtl.name = j.at("name").as_string();
    if (j.contains("albedo")) rgb(j.at("albedo"), mtl.m.albedo);
    if (j.contains("emission")) rgb(j.at("emission"), mtl.m.emission);
    if (j.contains("roughness")) mtl.m.roughness = (float)j.at("roughness").as_number();
    if (j.contains("metalness")) mtl.m.metallic = (float)j.at("metalness").as_number();
    if (j.contains("specular")) mtl.m.specular = (float)j.at("specular").as_number();
    if (j.contains("opacity")) mtl.m.opacity = (float)j.at("opacity").as_number();
    if (j.contains("transmission")) mtl.m.transmission = (float)j.at("transmission").as_number();
    if (j.contains("albedo_map")) mtl.albedo_map = j.at("albedo_map").as_string();
    if (j.contains("emission_map")) mtl.emission_map = j.at("emission_map").as_string();
    if (j.contains("roughness_map")) mtl.roughness_map = j.at("roughness_map").as_string();
    if (j.contains("metallic_map")) mtl.metallic_map = j.at("metallic_map").as_string();
    if (j.contains("normal_map")) mtl.normal_map = j.at("normal_map").as_string();
    if (j.contains("opacity_map")) mtl.opacity_map = j.at("opacity_map").as_string();
    if (j.contains("transmission_map")) mtl.transmission_map = j.at("transmission_map").as_string();
    if (j.contains("albedo_shader_id")) mtl.m.albedoShaderID = (int)j.at("albedo_shader_id").as_int64();
    // src/CommandManager.cpp:108-110: a DOUBLE square root (1.0 - anisotropic * 0.9 promotes), maxf is a<b?b:a on floats
    const float aspect = (float)sqrt(1.0 - mtl.m.anisotropic * 0.9);
    const float ax = mtl.m.roughness / aspect, ay = mtl.m.roughness * aspect;
    mtl.m.ax = 0.001f > ax ? 0.001f : ax;
    mtl.m.ay = 0.001f > ay ? 0.001f : ay;
    return mtl;
}

class CommandManager {
   public:
    TCPInterface* im = nullptr;
    er::RenderingManager rm;
    Scene scene;
    er::HostScene host_scene;  // what the running render borrows (kept alive until the next start)

    // one COMMAND message (its data messages are read here, src/main.cpp:129-172) -> the handler
    void execute(const Message& msg) {
        const std::string command_str = msg.get_string_data();
        InputCommand ic;
        try {
            ic = parse_command(command_str);
        } catch (const std::exception& e) {
            fprintf(stderr, "Input Command not recognized in: %s (%s)\n", command_str.c_str(), e.what());
            return;  // the reference logs and sends nothing (src/main.cpp:176-178, CommandManager.cpp:497-499)
        }
        if (ic.has("path")) {
            fprintf(stderr, "Not implemented yet: --path variants (src/CommandManager.cpp:116-236)\n");
            return;
        }
        if (ic.has("load_camera")) {
            const Message d = im->read_message();
            try {
                scene.camera = parse_camerajson(d.get_json_data(), scene.camera);
            } catch (const std::exception& e) {
                fprintf(stderr, "Invalid camera format: %s\n", e.what());
            }
            im->write_message(Message::OK());
        } else if (ic.has("load_texture")) {
            const Message meta = im->read_message(), data = im->read_message();
            try {
                scene.addTexture(parse_texturejson(meta.get_json_data(), data.get_float_data(), data.float_count()));
                scene.pair_textures();
            } catch (const std::exception& e) {
                fprintf(stderr, "Invalid texture format: %s\n", e.what());
            }
            im->write_message(Message::OK());
        } else if (ic.has("load_config")) {
            const Message d = im->read_message();
            try {
                const erj::Value j = d.get_json_data();
                er::RenderParameters rp;
                rp.width = (unsigned)j.at("x_res").as_int64();
                rp.height = (unsigned)j.at("y_res").as_int64();
                rp.sampleTarget = (unsigned)j.at("sample_target").as_int64();
                rp.denoise = j.at("denoise").as_bool();
                rp.device = j.at("device").as_string();
                rp.block_size = (unsigned)j.at("block_size").as_int64();
                rm.rend_pars = rp;
                scene.x_res = rp.width;
                scene.y_res = rp.height;
            } catch (const std::exception& e) {
                fprintf(stderr, "Invalid config format: %s\n", e.what());
            }
            im->write_message(Message::OK());
        } else if (ic.has("load_hdri")) {
            const Message meta = im->read_message(), data = im->read_message();
            try {
                Texture t = parse_texturejson(meta.get_json_data(), data.get_float_data(), data.float_count());
                if (ic.has("mirror_x")) t.mirror_x();
                if (ic.has("mirror_y")) t.mirror_y();
                t.pixel_shift(0.5f, 0);  // src/CommandManager.cpp:187
                scene.addHDRI(t);
            } catch (const std::exception& e) {
                fprintf(stderr, "Invalid config format: %s\n", e.what());
            }
            im->write_message(Message::OK());
        } else if (ic.has("load_brdf_material")) {
            const Message d = im->read_message();
            try {
                scene.addMaterial(parse_materialjson(d.get_json_data()));
                scene.pair_materials();
                scene.pair_textures();
            } catch (const std::exception& e) {
                fprintf(stderr, "Invalid material format: %s\n", e.what());
            }
            im->write_message(Message::OK());
        } else if (ic.has("load_object")) {
            const Message obj = im->read_message(), mtl = im->read_message();
            try {
                for (const LoadedObject& o : load_obj_text(obj.get_string_data(), mtl.get_string_data(), ic.has("recompute_normals")))
                    scene.addMeshObject(o.name, o.tris, o.matNames);
                scene.pair_materials();
            } catch (const std::exception& e) {
                fprintf(stderr, "Invalid .obj file: %s\n", e.what());
            }
            im->write_message(Message::OK());
        } else if (ic.has("start")) {
            try {
                scene.fill(host_scene);
                ErRenderOpts o{};
                rm.start_rendering(&host_scene, &o);
            } catch (const std::exception& e) {
                fprintf(stderr, "start failed: %s\n", e.what());
            }
            im->write_message(Message::OK());
        } else if (ic.has("get_info")) {  // src/CommandManager.cpp:278-300
            erj::Value j = erj::Value::make_object();
            unsigned samples = 0;
            try {
                samples = rm.get_render_info().samples;
            } catch (const std::exception& e) {
                fprintf(stderr, "Error copying render info: %s\n", e.what());
            }
            j.set("samples", erj::Value::of(samples));
            im->write_message(Message::json(j));
        } else if (ic.has("get_sycl_info")) {  // src/CommandManager.cpp:303-362: same keys, CUDA devices
            erj::Value j = erj::Value::make_object();
            erj::Value devices = erj::Value::make_array();
            try {
                for (const ErDeviceInfo& d : er::RenderingManager::devices()) {
                    erj::Value jd = erj::Value::make_object();
                    jd.set("name", erj::Value::of(d.name));
                    jd.set("platform", erj::Value::of(d.platform));
                    jd.set("memory", erj::Value::of((unsigned long long)d.memory_bytes));
                    jd.set("max_compute_units", erj::Value::of(d.max_compute_units));
                    jd.set("is_compatible", erj::Value::of(d.is_compatible != 0));
                    jd.set("online_compiler", erj::Value::of(false));  // kernels are compiled ahead of time for sm_100a
                    jd.set("type", erj::Value::of("gpu"));
                    devices.arr.push_back(jd);
                }
            } catch (const std::exception& e) {
                fprintf(stderr, "Invalid get_sycl_info: %s\n", e.what());
            }
            j.set("devices", devices);
            im->write_message(Message::json(j));
        } else if (ic.has("get_pass")) {  // src/CommandManager.cpp:250-276
            const size_t bytes = (size_t)rm.rend_pars.width * rm.rend_pars.height * 4 * sizeof(float);
            std::unique_ptr<float[]> pass;
            try {
                pass.reset(rm.get_pass(ic.get_pass));
            } catch (const std::exception& e) {
                fprintf(stderr, "get_pass failed: %s\n", e.what());
                pass.reset(new float[bytes / sizeof(float)]());
            }
            // rd.pars.denoise asks OpenImageDenoise for the beauty pass in the reference; OIDN is not part of this
            // build (network-fetched there, CMakeLists.txt:113-129): the raw pass goes out, alpha forced to 1 as there
            if (rm.rend_pars.denoise)
                for (size_t i = 3; i < bytes / sizeof(float); i += 4) pass[i] = 1;
            im->write_data(Message::DataFormat::FLOAT4, pass.get(), bytes);
        } else {
            fprintf(stderr, "Unrecognized input command: %s\n", command_str.c_str());
        }
    }
};

}  // namespace erh
