// PovRay-subset scene description front end.
//
// Same language and the same accept/expect recursive-descent behaviour as the reference's
// `SdlParser` (src/sdl_parser.rs): tokenizer separators (:282-308), dummy START token (:316),
// statement order (:434-478), quadric coefficient order (:659-661), modifier stack (:931-965),
// texture/finish/surface -> material mapping (:967-1100), #declare/#while/#end (:814-918),
// expression grammar (:1256-1402).  Runs once per scene on the host; nothing here is on the
// device path.
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <unordered_map>

#include "scene.hpp"

namespace raysnail {
namespace {

enum class Sym {
    Camera, Location, LookAt, Sphere, Box, Quadric, Light, Intersection, Difference, Object,
    Plus, Minus, Multiply, Divide, Equal, BlockOpen, BlockClose, VectorOpen, VectorClose,
    ParenOpen, ParenClose, Comma, Semicolon, Translate, Rotate, Scale, Texture, Pigment, Finish,
    Surface, Metallic, Reflection, Color, Rgb, Angle, Diffuse, Phong, PhongSize, Checker,
    Declare, While, End, Id, Eof, None
};

struct Token {
    std::string text;
    unsigned line;
};

struct Entity {
    enum Kind { Light, Camera, Hittable, Directive, Float, Vector, Invalid } kind = Invalid;
    LightData light;
    CameraData camera;
    HittablePtr hittable;
    double f = 0;
    Vec3 v;
    static Entity invalid() { return Entity(); }
    static Entity directive() { Entity e; e.kind = Directive; return e; }
};

const std::unordered_map<std::string, Sym>& symbol_map() {
    static const std::unordered_map<std::string, Sym> m = {
        {"camera", Sym::Camera}, {"look_at", Sym::LookAt}, {"location", Sym::Location},
        {"{", Sym::BlockOpen}, {"}", Sym::BlockClose}, {"intersection", Sym::Intersection},
        {"difference", Sym::Difference}, {"object", Sym::Object}, {"<", Sym::VectorOpen},
        {">", Sym::VectorClose}, {",", Sym::Comma}, {";", Sym::Semicolon}, {"sphere", Sym::Sphere},
        {"box", Sym::Box}, {"quadric", Sym::Quadric}, {"light", Sym::Light}, {"texture", Sym::Texture},
        {"pigment", Sym::Pigment}, {"finish", Sym::Finish}, {"surface", Sym::Surface},
        {"reflection", Sym::Reflection}, {"metallic", Sym::Metallic}, {"color", Sym::Color},
        {"rgb", Sym::Rgb}, {"checker", Sym::Checker}, {"angle", Sym::Angle}, {"diffuse", Sym::Diffuse},
        {"phong", Sym::Phong}, {"phong_size", Sym::PhongSize}, {"translate", Sym::Translate},
        {"rotate", Sym::Rotate}, {"scale", Sym::Scale}, {"+", Sym::Plus}, {"-", Sym::Minus},
        {"*", Sym::Multiply}, {"/", Sym::Divide}, {"(", Sym::ParenOpen}, {")", Sym::ParenClose},
        {"=", Sym::Equal}, {"#declare", Sym::Declare}, {"#while", Sym::While}, {"#end", Sym::End},
    };
    return m;
}

std::string trim(const std::string& s) {
    // Rust str::trim strips Unicode whitespace; ASCII whitespace is what scene files contain
    size_t a = 0, b = s.size();
    while (a < b && (s[a] == ' ' || s[a] == '\t' || s[a] == '\r' || s[a] == '\n' || s[a] == '\f' || s[a] == '\v')) ++a;
    while (b > a && (s[b - 1] == ' ' || s[b - 1] == '\t' || s[b - 1] == '\r' || s[b - 1] == '\n' || s[b - 1] == '\f' || s[b - 1] == '\v')) --b;
    return s.substr(a, b - a);
}

void push_non_empty(std::vector<Token>& v, const std::string& value, unsigned line) {
    std::string t = trim(value);
    if (!t.empty()) v.push_back({t, line});
}

bool is_sep(char c) {
    switch (c) {
        case ' ': case ',': case ';': case '(': case ')': case '<': case '>':
        case '{': case '}': case '+': case '-': case '*': case '/': case '\n':
            return true;
        default:
            return false;
    }
}

void tokenize(const std::string& line_in, unsigned line_no, std::vector<Token>& out) {
    std::string line = line_in;
    size_t c = line.find("//");
    if (c != std::string::npos) line = line.substr(0, c);
    std::string cur;
    for (char ch : line) {
        if (is_sep(ch)) {
            push_non_empty(out, cur, line_no);
            push_non_empty(out, std::string(1, ch), line_no);
            cur.clear();
        } else {
            cur.push_back(ch);
        }
    }
    if (!cur.empty()) push_non_empty(out, cur, line_no);
}

struct Input {
    size_t pos = 0;
    std::vector<Token> tokens;
    Sym symbol = Sym::None;
    std::unordered_map<std::string, Entity> declares;
    std::vector<size_t> loops;

    const std::string& current_text() const {
        return pos < tokens.size() ? tokens[pos].text : tokens.back().text;
    }
};

void nextsym(Input& in) {
    in.pos += 1;
    if (in.pos < in.tokens.size()) {
        auto it = symbol_map().find(in.tokens[in.pos].text);
        in.symbol = it == symbol_map().end() ? Sym::Id : it->second;
    } else {
        in.symbol = Sym::Eof;
    }
}

bool accept(Input& in, Sym s) {
    if (in.symbol == s) {
        nextsym(in);
        return true;
    }
    return false;
}
bool expect(Input& in, Sym s) { return accept(in, s); }

struct ParseFailure {
    std::string what;
};
[[noreturn]] void unwrap_none(const Input& in, const char* what) {
    // the reference calls Option::unwrap() here and would panic
    std::ostringstream ss;
    unsigned line = in.pos < in.tokens.size() ? in.tokens[in.pos].line : (unsigned)in.tokens.size();
    ss << what << " near line " << line << " ('" << in.current_text() << "')";
    throw ParseFailure{ss.str()};
}

bool parse_float(Input& in, double& out) {
    const std::string& t = in.current_text();
    if (t.empty()) return false;
    // accept what Rust's f64::from_str accepts for the tokens scene files contain
    char* end = nullptr;
    bool ok_chars = true;
    for (char ch : t)
        if (!((ch >= '0' && ch <= '9') || ch == '.' || ch == 'e' || ch == 'E' || ch == '+' || ch == '-')) ok_chars = false;
    if (!ok_chars) {
        if (t != "inf" && t != "infinity" && t != "nan" && t != "NaN") return false;
    }
    double v = std::strtod(t.c_str(), &end);
    if (end == t.c_str() || *end != '\0') return false;
    out = v;
    nextsym(in);
    return true;
}

bool parse_expression(Input& in, double& out);

bool parse_factor(Input& in, double& out) {
    if (accept(in, Sym::ParenOpen)) {
        double e = 0;
        bool ok = parse_expression(in, e);
        if (expect(in, Sym::ParenClose)) {
            out = e;
            return ok;
        }
        return false;
    }
    auto it = in.declares.find(in.current_text());
    if (it != in.declares.end() && it->second.kind == Entity::Float) {
        out = it->second.f;
        nextsym(in);
        return true;
    }
    return parse_float(in, out);
}

bool parse_term(Input& in, double& out) {
    double f;
    if (!parse_factor(in, f)) return false;
    for (;;) {
        double v;
        if (accept(in, Sym::Multiply)) {
            if (!parse_factor(in, v)) return false;
            f *= v;
        } else if (accept(in, Sym::Divide)) {
            if (!parse_factor(in, v)) return false;
            f /= v;
        } else {
            break;
        }
    }
    out = f;
    return true;
}

bool parse_expression(Input& in, double& out) {
    double e;
    if (accept(in, Sym::Minus)) {
        if (!parse_term(in, e)) return false;
        e = -e;
    } else {
        if (!parse_term(in, e)) return false;
    }
    for (;;) {
        double v;
        if (accept(in, Sym::Minus)) {
            if (!parse_term(in, v)) return false;
            e -= v;
        } else if (accept(in, Sym::Plus)) {
            if (!parse_term(in, v)) return false;
            e += v;
        } else {
            break;
        }
    }
    out = e;
    return true;
}

double expression_unwrap(Input& in) {
    double v;
    if (!parse_expression(in, v)) unwrap_none(in, "expected expression");
    return v;
}

bool parse_vector(Input& in, Vec3& out) {
    if (!expect(in, Sym::VectorOpen)) return false;
    double v1 = expression_unwrap(in);
    expect(in, Sym::Comma);
    double v2 = expression_unwrap(in);
    expect(in, Sym::Comma);
    double v3 = expression_unwrap(in);
    expect(in, Sym::VectorClose);
    out = Vec3(v1, v2, v3);
    return true;
}

Vec3 vector_unwrap(Input& in) {
    Vec3 v;
    if (!parse_vector(in, v)) unwrap_none(in, "expected vector");
    return v;
}

bool parse_color(Input& in, Color& out) {
    if (accept(in, Sym::Color)) {
        accept(in, Sym::Rgb);
        Vec3 v;
        if (parse_vector(in, v)) {
            out = Color::new64(v.x, v.y, v.z, 1.0);
            return true;
        }
    }
    return false;
}

bool parse_checker(Input& in, Color& c1, Color& c2) {
    if (accept(in, Sym::Checker)) {
        if (parse_color(in, c1)) {
            expect(in, Sym::Comma);
            if (parse_color(in, c2)) return true;
        }
    }
    return false;
}

TexturePtr parse_pigment(Input& in) {
    if (expect(in, Sym::Pigment)) {
        if (expect(in, Sym::BlockOpen)) {
            Color c, c2;
            if (parse_color(in, c)) {
                accept(in, Sym::Rgb);
                expect(in, Sym::BlockClose);
                return std::make_shared<SolidColor>(c);
            } else if (parse_checker(in, c, c2)) {
                expect(in, Sym::BlockClose);
                return std::make_shared<Checker>(std::make_shared<SolidColor>(c), std::make_shared<SolidColor>(c2), 2.0);
            }
        }
    }
    return nullptr;
}

CommonMaterialSettings make_settings(double phong_factor, double phong_exponent) {
    CommonMaterialSettings s;
    if (phong_factor > 0.0) {
        s.phong_factor = phong_factor * 4.0;
        s.phong_exponent = (int)(phong_exponent * 0.1);  // `as i32` truncates toward zero
    }
    return s;
}

double float_unwrap(Input& in) {
    double v;
    if (!parse_float(in, v)) unwrap_none(in, "expected float");
    return v;
}

MaterialPtr parse_finish(Input& in, TexturePtr texture) {
    if (expect(in, Sym::Finish)) {
        if (expect(in, Sym::BlockOpen)) {
            double phong = 0.0, phong_size = 40.0, reflection = 0.0;
            for (;;) {
                if (accept(in, Sym::Reflection)) reflection = float_unwrap(in);
                else if (accept(in, Sym::Phong)) phong = float_unwrap(in);
                else if (accept(in, Sym::PhongSize)) phong_size = float_unwrap(in);
                else break;
            }
            expect(in, Sym::BlockClose);
            if (reflection == 0.0) {
                auto l = std::make_shared<Lambertian>(texture);
                l->set(make_settings(phong, phong_size));
                return l;
            }
            auto l = std::make_shared<Lambertian>(texture);
            l->set(make_settings(phong, phong_size));
            auto m = std::make_shared<Metal>(texture);
            m->set(make_settings(phong, phong_size));
            return std::make_shared<MixedMaterial>(m, l, reflection);
        }
        // `finish` without `{`: falls out of the if/else chain to the Lambertian default
    } else if (expect(in, Sym::Surface)) {
        if (expect(in, Sym::BlockOpen)) {
            MaterialPtr material;
            if (accept(in, Sym::Metallic)) {
                if (accept(in, Sym::Diffuse)) {
                    double v = float_unwrap(in);
                    material = std::make_shared<DiffuseMetal>(v, texture);
                } else {
                    material = std::make_shared<Metal>(texture);
                }
            } else {
                material = std::make_shared<Lambertian>(texture);
            }
            expect(in, Sym::BlockClose);
            return material;
        }
    }
    return std::make_shared<Lambertian>(texture);
}

MaterialPtr parse_texture(Input& in) {
    if (accept(in, Sym::Texture)) {
        if (expect(in, Sym::BlockOpen)) {
            TexturePtr texture = parse_pigment(in);
            if (!texture) texture = std::make_shared<SolidColor>(Color(1.f, 1.f, 1.f, 1.f));
            MaterialPtr material = parse_finish(in, texture);
            expect(in, Sym::BlockClose);
            return material;
        }
    }
    return nullptr;
}

bool parse_translate(Input& in, Vec3& v) { return accept(in, Sym::Translate) && parse_vector(in, v); }
bool parse_rotate(Input& in, Vec3& v) { return accept(in, Sym::Rotate) && parse_vector(in, v); }
bool parse_scale(Input& in, Vec3& v) {
    if (accept(in, Sym::Scale)) {
        if (parse_vector(in, v)) return true;
        double f;
        if (parse_float(in, f)) {
            v = Vec3(f, f, f);
            return true;
        }
    }
    return false;
}

TransformStack parse_object_modifiers(Input& in) {
    const double PI = 3.14159265358979323846;
    TransformStack stack;
    for (;;) {
        Vec3 v;
        if (parse_translate(in, v)) {
            stack.push(Transform::translate(v));
        } else if (parse_rotate(in, v)) {
            if (v.x != 0.0) stack.push(Transform::rotate_by_x_axis(v.x * PI / 180.0));
            if (v.y != 0.0) stack.push(Transform::rotate_by_y_axis(v.y * PI / 180.0));
            if (v.z != 0.0) stack.push(Transform::rotate_by_z_axis(v.z * PI / 180.0));
        } else if (parse_scale(in, v)) {
            stack.push(Transform::scale(v));
        } else {
            break;
        }
    }
    return stack;
}

HittablePtr build_transform_facade(TransformStack stack, HittablePtr h) {
    if (stack.len() > 0) return std::make_shared<TfFacade>(std::move(h), std::move(stack));
    return h;
}

Entity hittable_entity(HittablePtr h) {
    Entity e;
    e.kind = Entity::Hittable;
    e.hittable = std::move(h);
    return e;
}

Entity parse_statement(Input& in);

Entity parse_camera(Input& in) {
    if (accept(in, Sym::Camera)) {
        if (expect(in, Sym::BlockOpen)) {
            CameraData cam;
            while (in.symbol != Sym::BlockClose) {
                if (in.symbol == Sym::Location) {
                    nextsym(in);
                    cam.location = vector_unwrap(in);
                } else if (in.symbol == Sym::LookAt) {
                    nextsym(in);
                    cam.look_at = vector_unwrap(in);
                } else if (in.symbol == Sym::Angle) {
                    nextsym(in);
                    cam.fov_angle = expression_unwrap(in);
                } else {
                    return Entity::invalid();
                }
            }
            nextsym(in);
            Entity e;
            e.kind = Entity::Camera;
            e.camera = cam;
            return e;
        }
    }
    return Entity::invalid();
}

Entity parse_light(Input& in) {
    if (accept(in, Sym::Light)) {
        if (expect(in, Sym::BlockOpen)) {
            LightData light;
            Vec3 loc;
            if (parse_vector(in, loc)) {
                expect(in, Sym::Comma);
                Color c;
                if (parse_color(in, c)) {
                    expect(in, Sym::BlockClose);
                    light.location = loc;
                    light.color = c;
                } else {
                    return Entity::invalid();
                }
            } else {
                return Entity::invalid();
            }
            Entity e;
            e.kind = Entity::Light;
            e.light = light;
            return e;
        }
    }
    return Entity::invalid();
}

Entity parse_sphere(Input& in) {
    if (accept(in, Sym::Sphere)) {
        if (expect(in, Sym::BlockOpen)) {
            Vec3 v = vector_unwrap(in);
            expect(in, Sym::Comma);
            double r = expression_unwrap(in);
            MaterialPtr material = parse_texture(in);
            TransformStack stack = parse_object_modifiers(in);
            auto sphere = std::make_shared<Sphere>(v, r, material);
            expect(in, Sym::BlockClose);
            return hittable_entity(build_transform_facade(std::move(stack), sphere));
        }
    }
    return Entity::invalid();
}

Entity parse_box(Input& in) {
    if (accept(in, Sym::Box)) {
        if (expect(in, Sym::BlockOpen)) {
            Vec3 v1 = vector_unwrap(in);
            expect(in, Sym::Comma);
            Vec3 v2 = vector_unwrap(in);
            MaterialPtr material = parse_texture(in);
            TransformStack stack = parse_object_modifiers(in);
            auto gbox = std::make_shared<Box>(v1, v2, material);
            expect(in, Sym::BlockClose);
            return hittable_entity(build_transform_facade(std::move(stack), gbox));
        }
    }
    return Entity::invalid();
}

Entity parse_quadric(Input& in) {
    if (accept(in, Sym::Quadric)) {
        if (expect(in, Sym::BlockOpen)) {
            Vec3 v1 = vector_unwrap(in);
            expect(in, Sym::Comma);
            Vec3 v2 = vector_unwrap(in);
            expect(in, Sym::Comma);
            Vec3 v3 = vector_unwrap(in);
            expect(in, Sym::Comma);
            double j = expression_unwrap(in);
            MaterialPtr material = parse_texture(in);
            TransformStack stack = parse_object_modifiers(in);
            auto quadric = std::make_shared<Quadric>(v1.x, v2.x, v2.y, v3.x, v1.y, v2.z, v3.y, v1.z, v3.z, j, material);
            expect(in, Sym::BlockClose);
            return hittable_entity(build_transform_facade(std::move(stack), quadric));
        }
    }
    return Entity::invalid();
}

Entity parse_object(Input& in) {
    if (accept(in, Sym::Object)) {
        if (expect(in, Sym::BlockOpen)) {
            std::string ident = in.current_text();
            nextsym(in);
            TransformStack stack = parse_object_modifiers(in);
            expect(in, Sym::BlockClose);
            auto it = in.declares.find(ident);
            if (it == in.declares.end()) unwrap_none(in, "undeclared identifier");
            if (it->second.kind == Entity::Hittable)
                return hittable_entity(build_transform_facade(std::move(stack), it->second.hittable));
        }
    }
    return Entity::invalid();
}

template <class CSG>
Entity parse_csg(Input& in, Sym keyword) {
    if (accept(in, keyword)) {
        if (expect(in, Sym::BlockOpen)) {
            Entity a = parse_statement(in);
            if (a.kind == Entity::Hittable) {
                Entity b = parse_statement(in);
                if (b.kind == Entity::Hittable) {
                    MaterialPtr material = parse_texture(in);
                    TransformStack stack = parse_object_modifiers(in);
                    auto csg = std::make_shared<CSG>(a.hittable, b.hittable, material);
                    expect(in, Sym::BlockClose);
                    return hittable_entity(build_transform_facade(std::move(stack), csg));
                }
            }
        }
    }
    return Entity::invalid();
}

Entity parse_declare(Input& in) {
    if (accept(in, Sym::Declare)) {
        std::string ident = in.current_text();
        nextsym(in);
        if (expect(in, Sym::Equal)) {
            double f;
            Vec3 v;
            if (parse_expression(in, f)) {
                nextsym(in);  // consumes the ';' (sdl_parser.rs:826-828)
                expect(in, Sym::Semicolon);
                Entity e;
                e.kind = Entity::Float;
                e.f = f;
                in.declares[ident] = e;
                return Entity::directive();
            } else if (parse_vector(in, v)) {
                expect(in, Sym::Semicolon);
                Entity e;
                e.kind = Entity::Vector;
                e.v = v;
                in.declares[ident] = e;
                return Entity::directive();
            } else {
                in.declares[ident] = parse_statement(in);
                return Entity::directive();
            }
        }
    }
    return Entity::invalid();
}

void fast_forward_to_end(Input& in) {
    while (in.symbol != Sym::End) {
        if (in.symbol == Sym::Eof) throw ParseFailure{"#while without #end"};
        nextsym(in);
    }
    nextsym(in);
}

Entity parse_while(Input& in) {
    size_t loop_start = in.pos;
    if (accept(in, Sym::While)) {
        if (expect(in, Sym::ParenOpen)) {
            double v1, v2;
            if (parse_expression(in, v1)) {
                nextsym(in);  // skips the '<'
                expect(in, Sym::VectorOpen);
                if (parse_expression(in, v2)) {
                    nextsym(in);  // skips the ')'
                    if (v1 < v2) in.loops.push_back(loop_start);
                    else fast_forward_to_end(in);
                    return Entity::directive();
                }
            }
        }
    }
    return Entity::invalid();
}

Entity parse_end(Input& in) {
    if (accept(in, Sym::End)) {
        if (in.loops.empty()) throw ParseFailure{"#end without #while"};
        size_t loop_start = in.loops.back();
        in.loops.pop_back();
        in.pos = loop_start - 1;
        nextsym(in);
        return Entity::directive();
    }
    return Entity::invalid();
}

Entity parse_statement(Input& in) {
    Entity e;
    if ((e = parse_camera(in)).kind != Entity::Invalid) return e;
    if ((e = parse_light(in)).kind != Entity::Invalid) return e;
    if ((e = parse_sphere(in)).kind != Entity::Invalid) return e;
    if ((e = parse_box(in)).kind != Entity::Invalid) return e;
    if ((e = parse_quadric(in)).kind != Entity::Invalid) return e;
    if ((e = parse_object(in)).kind != Entity::Invalid) return e;
    if ((e = parse_csg<Difference>(in, Sym::Difference)).kind != Entity::Invalid) return e;
    if ((e = parse_csg<Intersection>(in, Sym::Intersection)).kind != Entity::Invalid) return e;
    if ((e = parse_declare(in)).kind != Entity::Invalid) return e;
    if ((e = parse_while(in)).kind != Entity::Invalid) return e;
    if ((e = parse_end(in)).kind != Entity::Invalid) return e;
    return Entity::invalid();
}

bool parse_tokens(Input& in, SceneData& scene, std::string& error) {
    try {
        nextsym(in);
        while (in.pos < in.tokens.size()) {
            Entity e = parse_statement(in);
            switch (e.kind) {
                case Entity::Hittable: scene.hittables.add_ref(e.hittable); break;
                case Entity::Light: scene.lights.push_back(e.light); break;
                case Entity::Camera: scene.camera = e.camera; scene.has_camera = true; break;
                case Entity::Directive: case Entity::Float: case Entity::Vector: break;
                case Entity::Invalid: {
                    std::ostringstream ss;
                    unsigned line = in.pos < in.tokens.size() ? in.tokens[in.pos].line : 0;
                    ss << "Parse error at line " << line << " near '" << in.current_text() << "'";
                    error = ss.str();
                    return false;
                }
            }
        }
    } catch (const ParseFailure& f) {
        error = "Parse error: " + f.what;
        return false;
    } catch (const std::exception& ex) {
        error = std::string("Parse error: ") + ex.what();
        return false;
    }
    return true;
}

}  // namespace

bool SdlParser::parse_text(const std::string& text, SceneData& out, std::string& error) {
    Input in;
    in.tokens.push_back({"START", 0});
    std::istringstream ss(text);
    std::string line;
    unsigned line_no = 1;
    while (std::getline(ss, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        tokenize(line, line_no, in.tokens);
        ++line_no;
    }
    return parse_tokens(in, out, error);
}

bool SdlParser::parse(const std::string& filename, SceneData& out, std::string& error) {
    std::ifstream f(filename);
    if (!f) {
        error = "cannot read " + filename;
        return false;
    }
    std::stringstream buf;
    buf << f.rdbuf();
    return parse_text(buf.str(), out, error);
}

// src/bin/raysnail.rs:340-377
std::unique_ptr<World> world_from_scene_data(SceneData& sd, CameraBuilder& cam_out) {
    cam_out = CameraBuilder();
    cam_out.look_from(sd.camera.location).look_at(sd.camera.look_at).fov(sd.camera.fov_angle).aperture(0.01).focus(10.0);
    HittableList lights;
    for (const LightData& l : sd.lights) {
        auto mat = std::make_shared<DiffuseLight>(std::make_shared<SolidColor>(l.color));
        mat->multiplier(1.7);
        auto rs = std::make_shared<Sphere>(l.location, 12.0, mat);
        lights.add_ref(rs);
        sd.hittables.add_ref(rs);
    }
    return std::make_unique<World>(sd.hittables, lights, Color(0.3f, 0.4f, 0.5f, 1.0f), Color(0.7f, 0.89f, 1.0f, 1.0f),
                                   TimeRange{0.0, 0.0});
}

}  // namespace raysnail
