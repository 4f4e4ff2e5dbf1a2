// visionaray_b200/obj_loader.inl -- bodies of load_obj() (see obj_loader.h for the rules and what they follow).

namespace visionaray_b200
{
namespace detail
{

// binary PNM (P6 RGB, P5 gray, maxval 255) -> RGBA8 with alpha 1 (obj_loader.cpp:400-424: PF_RGB8 / PF_R8 branches)
inline bool obj_load_pnm(std::string const& path, texture2d& tex)
{
    std::ifstream f(path.c_str(), std::ios::binary);
    if (!f.good()) return false;
    std::string magic;
    f >> magic;
    if (magic != "P6" && magic != "P5") return false;
    long vals[3]; int got = 0;
    while (got < 3 && f.good())
    {
        int c = f.peek();
        if (c == '#') { std::string skip; std::getline(f, skip); continue; }
        if (c == ' ' || c == '\t' || c == '\n' || c == '\r') { f.get(); continue; }
        if (!(f >> vals[got])) return false;
        ++got;
    }
    if (got < 3 || vals[0] <= 0 || vals[1] <= 0 || vals[2] != 255) return false;
    f.get();                                              // the single whitespace after maxval
    const size_t w = size_t(vals[0]), h = size_t(vals[1]), ch = magic == "P6" ? 3 : 1;
    std::vector<unsigned char> raw(w * h * ch);
    f.read(reinterpret_cast<char*>(raw.data()), std::streamsize(raw.size()));
    if (size_t(f.gcount()) != raw.size()) return false;
    tex.width = w; tex.height = h; tex.texels.resize(w * h);
    for (size_t i = 0; i < w * h; ++i)
    {
        unorm8x4 t;
        if (ch == 3) { t.x = raw[3 * i]; t.y = raw[3 * i + 1]; t.z = raw[3 * i + 2]; }
        else { t.x = t.y = t.z = raw[i]; }
        t.w = 255;
        tex.texels[i] = t;
    }
    return true;
}

// parse_mtl, obj_loader.cpp:193-241: newmtl / Ka / Kd / Ke / Ks / Ns / map_Kd; a name seen twice keeps its first entry
// as the current one; anything else is skipped
inline void obj_parse_mtl(std::string const& filename, std::map<std::string, obj_mtl>& matlib)
{
    std::ifstream f(filename.c_str(), std::ios::binary);
    std::string line, key, rest;
    std::map<std::string, obj_mtl>::iterator cur = matlib.end();
    while (std::getline(f, line))
    {
        obj_split(line, key, rest);
        if (key == "newmtl") { cur = matlib.insert(std::make_pair(rest, obj_mtl())).first; continue; }
        if (cur == matlib.end()) continue;
        float v[3];
        if (key == "Ka" && obj_floats(rest, v, 3) == 3) cur->second.ka = vec3(v[0], v[1], v[2]);
        else if (key == "Kd" && obj_floats(rest, v, 3) == 3) cur->second.kd = vec3(v[0], v[1], v[2]);
        else if (key == "Ke" && obj_floats(rest, v, 3) == 3) cur->second.ke = vec3(v[0], v[1], v[2]);
        else if (key == "Ks" && obj_floats(rest, v, 3) == 3) cur->second.ks = vec3(v[0], v[1], v[2]);
        else if (key == "Ns" && obj_floats(rest, v, 1) == 1) cur->second.ns = v[0];
        else if (key == "map_Kd") cur->second.map_kd = rest;
    }
}

// store_triangle, obj_loader.cpp:65-88
inline bool obj_store_triangle(model& result, std::vector<vec3> const& vertices, int i1, int i2, int i3)
{
    model::triangle_type tri;
    tri.v1 = vertices[size_t(i1)];
    tri.e1 = vertices[size_t(i2)] - tri.v1;
    tri.e2 = vertices[size_t(i3)] - tri.v1;
    if (length(cross(tri.e1, tri.e2)) == 0.0f)
    {
        std::ostringstream w;
        w << "rejecting degenerate triangle: zero-based indices: (" << i1 << ' ' << i2 << ' ' << i3 << ")";
        result.warnings.push_back(w.str());
        return false;
    }
    tri.prim_id = unsigned(result.primitives.size());
    tri.geom_id = result.materials.empty() ? 0u : unsigned(result.materials.size() - 1);
    result.primitives.push_back(tri);
    return true;
}

// store_faces, obj_loader.cpp:95-146
inline void obj_store_faces(model& result, std::vector<vec3> const& vertices, std::vector<vec2> const& tex_coords,
                            std::vector<vec3> const& normals, std::vector<obj_face_index> const& faces)
{
    const int nv = int(vertices.size()), nt = int(tex_coords.size()), nn = int(normals.size());
    auto ok = [](int i, int n) { return i >= 0 && i < n; };
    const int i1 = obj_remap_index(faces[0].v, nv);
    for (size_t last = 2; last != faces.size(); ++last)
    {
        const int i2 = obj_remap_index(faces[last - 1].v, nv), i3 = obj_remap_index(faces[last].v, nv);
        if (!ok(i1, nv) || !ok(i2, nv) || !ok(i3, nv))          // the reference indexes out of bounds here; refused instead
        {
            result.warnings.push_back("face refers to a vertex that does not exist");
            continue;
        }
        if (!obj_store_triangle(result, vertices, i1, i2, i3)) continue;
        if (faces[0].has_t && faces[last - 1].has_t && faces[last].has_t)
        {
            const int t1 = obj_remap_index(faces[0].t, nt), t2 = obj_remap_index(faces[last - 1].t, nt), t3 = obj_remap_index(faces[last].t, nt);
            const vec2 zero{ 0.0f, 0.0f };
            result.tex_coords.push_back(ok(t1, nt) ? tex_coords[size_t(t1)] : zero);
            result.tex_coords.push_back(ok(t2, nt) ? tex_coords[size_t(t2)] : zero);
            result.tex_coords.push_back(ok(t3, nt) ? tex_coords[size_t(t3)] : zero);
        }
        if (faces[0].has_n && faces[last - 1].has_n && faces[last].has_n)
        {
            const int n1 = obj_remap_index(faces[0].n, nn), n2 = obj_remap_index(faces[last - 1].n, nn), n3 = obj_remap_index(faces[last].n, nn);
            const vec3 up(0.0f, 0.0f, 1.0f);
            result.shading_normals.push_back(ok(n1, nn) ? normals[size_t(n1)] : up);
            result.shading_normals.push_back(ok(n2, nn) ? normals[size_t(n2)] : up);
            result.shading_normals.push_back(ok(n3, nn) ? normals[size_t(n3)] : up);
        }
    }
}

} // namespace detail

// load_obj(filename, model&) -- src/common/obj_loader.cpp:299-533.  Throws std::runtime_error when the file cannot be
// opened (the reference's mapped_file_source throws there too).
inline void load_obj(std::string const& filename, model& mod)
{
    using namespace detail;
    std::ifstream file(filename.c_str(), std::ios::binary);
    if (!file.good()) throw std::runtime_error("load_obj: cannot open " + filename);

    std::map<std::string, obj_mtl> matlib;
    size_t geom_id = 0;
    std::vector<vec3> vertices, normals;
    std::vector<vec2> tex_coords;
    std::vector<obj_face_index> faces;
    std::string line, key, rest;
    const std::string dir = obj_dirname(filename);

    while (std::getline(file, line))
    {
        obj_split(line, key, rest);
        if (key.empty() || key[0] == '#') continue;
        float v[6];
        if (key == "mtllib")
        {
            const std::string mtl_path = dir + "/" + rest;
            if (obj_file_exists(mtl_path)) obj_parse_mtl(mtl_path, matlib);
            else mod.warnings.push_back("file does not exist: " + mtl_path);
        }
        else if (key == "usemtl")
        {
            std::map<std::string, obj_mtl>::const_iterator it = matlib.find(rest);
            if (it != matlib.end())
            {
                obj_mtl const& m = it->second;
                plastic<float> mat;                                  // add_material(plastic), obj_loader.cpp:247-262
                mat.set_ca(m.ka); mat.set_cd(m.kd); mat.set_cs(m.ks);
                mat.set_ka(1.0f); mat.set_kd(1.0f); mat.set_ks(1.0f);
                mat.set_specular_exp(m.ns);
                mod.materials.push_back(mat);
                if (!m.map_kd.empty())
                {
                    std::string tex_filename = m.map_kd[0] == '/' ? m.map_kd : std::string();
                    if (tex_filename.empty() || !obj_file_exists(tex_filename))
                    {
                        tex_filename = dir + "/" + m.map_kd;             // relative to the obj file, '\\' -> '/'
                        for (char& c : tex_filename) if (c == '\\') c = '/';
                    }
                    if (obj_file_exists(tex_filename))
                    {
                        std::map<std::string, texture2d>::iterator t = mod.texture_map.find(m.map_kd);
                        if (t == mod.texture_map.end())
                        {
                            texture2d tex;
                            if (obj_load_pnm(tex_filename, tex)) t = mod.texture_map.insert(std::make_pair(m.map_kd, std::move(tex))).first;
                            else mod.warnings.push_back("cannot load texture from file (binary PNM only): " + tex_filename);
                        }
                        if (t != mod.texture_map.end()) mod.textures.push_back(t->second.ref());
                    }
                    else mod.warnings.push_back("file does not exist: " + tex_filename);
                }
                if (mod.textures.size() < mod.materials.size()) mod.textures.push_back(texture_ref2d(0, 0));   // empty dummy
            }
            else mod.warnings.push_back("material not present in mtllib: " + rest);
            geom_id = mod.materials.empty() ? 0 : mod.materials.size() - 1;
        }
        else if (key == "v")  { if (obj_floats(rest, v, 6) >= 3) vertices.push_back(vec3(v[0], v[1], v[2])); }   // w / RGB extension ignored
        else if (key == "vt") { if (obj_floats(rest, v, 3) >= 2) tex_coords.push_back(vec2{ v[0], v[1] }); }
        else if (key == "vn") { if (obj_floats(rest, v, 3) == 3) normals.push_back(vec3(v[0], v[1], v[2])); }
        else if (key == "f")
        {
            faces.clear();
            char const* p = rest.c_str();
            obj_face_index fi;
            while (obj_face_token(p, fi)) faces.push_back(fi);
            if (faces.size() >= 3) obj_store_faces(mod, vertices, tex_coords, normals, faces);
        }
    }
    // texture refs point into texture_map's nodes, which std::map never moves: still valid here

    for (model::triangle_type const& tri : mod.primitives) mod.geometric_normals.push_back(normalize(cross(tri.e1, tri.e2)));
    for (size_t i = mod.tex_coords.size() / 3; i < mod.primitives.size(); ++i)
        for (int k = 0; k < 3; ++k) mod.tex_coords.push_back(vec2{ 0.0f, 0.0f });
    for (size_t i = mod.materials.size(); i <= geom_id; ++i) mod.materials.push_back(make_default_material());
    for (size_t i = mod.textures.size(); i <= geom_id; ++i) mod.textures.push_back(texture_ref2d(0, 0));
    mod.bbox.invalidate();
    for (model::triangle_type const& tri : mod.primitives)
    {
        mod.bbox.insert(tri.v1); mod.bbox.insert(tri.v1 + tri.e1); mod.bbox.insert(tri.v1 + tri.e2);
    }
}

} // namespace visionaray_b200
