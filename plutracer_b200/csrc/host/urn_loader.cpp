// urn_loader.cpp — plu_scene_load_urn: the ".urn" scene description language and the scene builder,
// producing the flat tables directly (no object graph).
//
// Replaces, for the render path's input side, urn::token_stream / urn::value / urn::eval_context
// (reference inc/urn.h, src/urn.cpp, docs/lang.md) and plu::scenes::scene (inc/scene.h:229-298).
// Written from the language description and the observable behaviour of the reference loader, which
// tests/test_urn_loader.py pins byte-for-byte: the .pluscene this loader writes for a scene file must
// equal the one the reference process dumps for the same file (tests/golden/make_golden.py).
// Behaviours kept on purpose: a blank line ends parsing (urn.h:106-108), numbers have no exponent
// form (urn.h:63-68), `up:` is never read, `/res WxH` and `/smp N` override the file (scene.h:232-256).
#include "bvh_build.h"
#include "scene_store.h"
#include <array>
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <list>
#include <map>
#include <memory>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace {

struct parse_error : std::runtime_error { using std::runtime_error::runtime_error; };

// ---- tokens -------------------------------------------------------------------------------------------
struct token { enum kind { id, num, str, ctrl } type; std::string value; };

// Line-oriented cursor with the reference's end-of-input rule: input is exhausted as soon as the cursor
// stands on an empty line.
class token_stream {
	std::vector<std::string> lines;
	size_t ln = 0, col = 0;
	static bool term(char c) { return std::isspace((unsigned char)c) || c == '(' || c == ')' || c == '[' || c == ']'; }
	void advance(size_t n) {
		col += n;   // on an empty line the cursor stays put: more() is then false for good (urn.h:40-45)
		if (!lines[ln].empty() && col > lines[ln].size() - 1) { col = 0; ln++; }
	}
	void skip_ws() {
		while (more() && (std::isspace((unsigned char)lines[ln][col]) || std::iscntrl((unsigned char)lines[ln][col]))) advance(1);
	}
	// token at the cursor (+ how many characters to consume)
	std::pair<token, size_t> get() const {
		if (ln >= lines.size()) throw parse_error("unexpected end of input");
		const std::string* lb = &lines[ln];
		size_t i = col;
		while (i < lb->size() && std::isspace((unsigned char)(*lb)[i])) i++;
		if (i >= lb->size()) {
			if (ln + 1 >= lines.size()) throw parse_error("unexpected end of input");
			lb = &lines[ln + 1]; i = 0;
			while (i < lb->size() && std::isspace((unsigned char)(*lb)[i])) i++;
			if (i >= lb->size()) throw parse_error("unexpected end of input (blank line)");
		}
		char c = (*lb)[i];
		if (c == '(' || c == ')' || c == '[' || c == ']') return {token{token::ctrl, std::string(1, c)}, 1};
		if (c == '-' || std::isdigit((unsigned char)c)) {
			std::string v;
			do { v += (*lb)[i++]; } while (i < lb->size() && ((*lb)[i] == '.' || std::isdigit((unsigned char)(*lb)[i])));
			return {token{token::num, v}, v.size()};
		}
		if (c == '"') {
			i++;
			std::string v;
			while (i < lb->size() && (*lb)[i] != '"') v += (*lb)[i++];
			if (i >= lb->size()) throw parse_error("string literal didn't close");
			return {token{token::str, v}, v.size() + 2};
		}
		std::string v;
		while (i < lb->size() && !term((*lb)[i])) v += (*lb)[i++];
		return {token{token::id, v}, v.size()};
	}
public:
	explicit token_stream(std::istream& in) {
		std::string s;
		while (in.good()) { std::getline(in, s); lines.push_back(s); }
	}
	bool more() const { return ln < lines.size() && col < lines[ln].size(); }
	token next() { skip_ws(); auto x = get(); advance(x.second); return x.first; }
	token peek() const { return get().first; }
};

// ---- values -------------------------------------------------------------------------------------------
// texture2d::texture2d(bmp_filename), texture.cpp:4-65, restated with its quirks: 24-bit uncompressed only; the
// pixel loop runs i = imageSize, imageSize-3, .., 3 over data[i], data[i+1], data[i+2], so texel j is the BMP
// pixel imageSize/3 - j (the image comes out rotated by 180 degrees and shifted by one pixel), row padding bytes
// are taken for pixels, the first pixel of the file is never used and texel 0 is read one pixel PAST the buffer
// (heap garbage in the reference; 0 here).  Appends width*height texels (what texture2d::texel can address).
static bool load_bmp_texels(const std::string& path, std::vector<float>& out, int32_t& width, int32_t& height) {
	FILE* f = fopen(path.c_str(), "rb");
	if (!f) return false;
	unsigned char header[54];
	if (fread(header, 1, 54, f) != 54 || header[0] != 'B' || header[1] != 'M') { fclose(f); throw parse_error("invalid BMP file '" + path + "'"); }
	auto i32 = [&](int at) { int32_t v; std::memcpy(&v, header + at, 4); return v; };
	if (i32(0x1E) != 0 || i32(0x1C) != 24) { fclose(f); throw parse_error("Not a correct BMP file '" + path + "' (24 bpp, uncompressed)"); }
	uint32_t image_size = (uint32_t)i32(0x22);
	width = i32(0x12); height = i32(0x16);
	if (width <= 0 || height <= 0 || (int64_t)width * height > (int64_t)1 << 28) { fclose(f); throw parse_error("bad BMP size in '" + path + "'"); }
	if (image_size == 0) image_size = (uint32_t)width * (uint32_t)height * 3u;
	std::vector<unsigned char> data((size_t)image_size + 3, 0);   // +3: the out-of-bounds pixel reads as 0
	size_t got = fread(data.data(), 1, image_size, f);   // the reference reads from the current position (54), ignoring dataPos
	(void)got;
	fclose(f);
	size_t n_px = image_size / 3, want = (size_t)width * (size_t)height, j = 0;
	size_t base = out.size();
	out.resize(base + want * 3, 0.f);
	for (int64_t i = (int64_t)image_size; i > 0 && j < want; i -= 3, ++j) {
		out[base + 3 * j + 0] = (float)data[(size_t)i + 2] / 255.f;   // d.r
		out[base + 3 * j + 1] = (float)data[(size_t)i + 1] / 255.f;   // d.g
		out[base + 3 * j + 2] = (float)data[(size_t)i] / 255.f;       // d.b
	}
	(void)n_px;
	return true;
}

struct value;
using value_list = std::vector<value>;
struct func_t;

struct value {
	enum kind { Null, Int, Float, String, Var, Id, Def, Val, Block, Group, Func, Native } type = Null;
	int64_t i = 0;
	double f = 0;
	std::string s;
	value_list vs;
	std::shared_ptr<func_t> fn;
	std::shared_ptr<std::function<value()>> native;

	value() {}
	static value integer(int64_t v) { value r; r.type = Int; r.i = v; return r; }
	static value real(double v) { value r; r.type = Float; r.f = v; return r; }
	static value block(value_list v) { value r; r.type = Block; r.vs = std::move(v); return r; }   // (callers hand their local list over)
	static value nativefn(std::function<value()> f) { value r; r.type = Native; r.native = std::make_shared<std::function<value()>>(std::move(f)); return r; }

	bool is_null() const { return type == Null; }
	double num() const {
		if (type == Int) return (double)i;
		if (type == Float) return f;
		throw parse_error("expected value of type Int or Float");
	}
	int64_t integer_() const { if (type != Int) throw parse_error("expected value of type Int"); return i; }
	const std::string& var() const { if (type != Var) throw parse_error("expected value of type Var"); return s; }
	const std::string& ident() const { if (type != Id) throw parse_error("expected value of type Id"); return s; }
	const std::string& str() const { if (type != String) throw parse_error("expected value of type String"); return s; }
	const value_list& items() const { if (type != Block) throw parse_error("expected value of type Block"); return vs; }
	const value& at(size_t k) const {
		if (!(type == Block || type == Group)) throw parse_error("expected value of type Block or Group");
		if (k >= vs.size()) throw parse_error("block has too few values");
		return vs[k];
	}
	// `[a: 3 b: 7]`.named("a") -> 3
	value named(const std::string& nm) const {
		if (type != Block) throw parse_error("expected value of type Block");
		for (const auto& v : vs) if (v.type == Def && v.s == nm) return v.vs[0];
		return value();
	}
	bool has_named(const std::string& nm) const {
		if (type != Block) return false;
		for (const auto& v : vs) if (v.type == Def && v.s == nm) return true;
		return false;
	}
};
struct func_t { std::vector<std::string> argnames; value body; };

value parse(token_stream& ts) {
	token t = ts.next();
	value v;
	if (t.type == token::num) {
		if (t.value.find('.') != std::string::npos) { v.type = value::Float; v.f = std::atof(t.value.c_str()); }
		else { v.type = value::Int; v.i = std::atoll(t.value.c_str()); }
	} else if (t.type == token::str) {
		v.type = value::String; v.s = t.value;
	} else if (t.type == token::id) {
		if (t.value.empty()) throw parse_error("empty identifier");
		if (t.value[0] == ':') { v.type = value::Val; v.s = t.value.substr(1); }
		else if (t.value[0] == '\'') { v.type = value::Id; v.s = t.value.substr(1); }
		else if (t.value.back() == ':') { v.type = value::Def; v.s = t.value.substr(0, t.value.size() - 1); v.vs.push_back(parse(ts)); }
		else { v.type = value::Var; v.s = t.value; }
	} else {
		std::string close;
		if (t.value == "(") { close = ")"; v.type = value::Group; }
		else if (t.value == "[") { close = "]"; v.type = value::Block; }
		else throw parse_error("unexpected '" + t.value + "'");
		while (ts.more()) {
			token nt = ts.peek();
			if (nt.type == token::ctrl && nt.value == close) { ts.next(); break; }
			v.vs.push_back(parse(ts));
		}
	}
	return v;
}

// ---- evaluation (docs/lang.md "evaluation", src/urn.cpp:196-395) ---------------------------------------
class eval_context {
	std::list<std::map<std::string, value>> scopes;
	std::set<std::string> binary_ops;
	void push() { scopes.emplace_front(); }
	void pop() { scopes.pop_front(); }
	void bind(const std::string& n, const value& v) { scopes.front()[n] = v; }
	value lookup(const std::string& n) const {
		for (const auto& sc : scopes) { auto it = sc.find(n); if (it != sc.end()) return it->second; }
		return value();
	}
	void def_native(const std::string& name, std::vector<std::string> args, std::function<value()> body) {
		value f; f.type = value::Func; f.fn = std::make_shared<func_t>();
		f.fn->argnames = std::move(args); f.fn->body = value::nativefn(std::move(body));
		bind(name, f);
	}
	static value format(const value_list& fmt, const value_list& vals) {   // block-format
		value_list out;
		for (const auto& v : fmt) {
			if (v.type == value::Val) {
				char* end = nullptr;
				long idx = std::strtol(v.s.c_str(), &end, 10);
				if (end != v.s.c_str()) { if (idx < 0 || (size_t)idx >= vals.size()) throw parse_error("block-format index out of range"); out.push_back(vals[(size_t)idx]); }
				else out.push_back(v);
			} else if (v.type == value::Block) out.push_back(format(v.vs, vals));
			else out.push_back(v);
		}
		return value::block(std::move(out));
	}
	// one expression starting at Vs[i]: a function call with its arguments, or a value followed by
	// any number of `binary-op value` pairs (left to right, no precedence)
	value expr(const value_list& Vs, size_t& i) {
		value fv = eval1(Vs[i]);
		if (Vs[i].type != value::Val && fv.type == value::Func) {
			func_t F = *fv.fn;
			push();
			i++;
			for (size_t j = 0; i < Vs.size() && j < F.argnames.size(); ++i, ++j) {
				value a = eval1(Vs[i]);
				bind(F.argnames[j], a);
			}
			value rv = eval(F.body);
			pop();
			return rv;
		}
		while (i + 1 < Vs.size() && Vs[i + 1].type == value::Var && binary_ops.count(Vs[i + 1].s)) {
			func_t F = *lookup(Vs[i + 1].s).fn;
			i += 2;
			if (i >= Vs.size()) throw parse_error("binary operator without a right operand");
			push();   // the right operand is resolved inside the operator's scope, after `a` is bound (urn.cpp:259-262)
			bind(F.argnames[0], fv);
			value rhs = eval1(Vs[i]);
			bind(F.argnames[1], rhs);
			fv = eval(F.body);
			pop();
		}
		i++;
		return fv;
	}
public:
	eval_context() {
		push();
		binary_ops.insert("+");
		def_native("+", {"a", "b"}, [this]() -> value {
			value a = lookup("a"), b = lookup("b");
			if (a.type == value::Int && b.type == value::Int) return value::integer(a.i + b.i);
			if (a.type == value::Int && b.type == value::Float) return value::real((double)a.i + b.f);
			if (a.type == value::Float && b.type == value::Int) return value::real(a.f + (double)b.i);
			if (a.type == value::Float && b.type == value::Float) return value::real(a.f + b.f);
			if (a.type == value::Block && b.type == value::Block) { value_list v = a.vs; v.insert(v.end(), b.vs.begin(), b.vs.end()); return value::block(std::move(v)); }
			return value();
		});
		def_native("do", {"v"}, [this]() { return eval(lookup("v")); });
		def_native("reduce", {"v"}, [this]() { return reduce(lookup("v")); });
		def_native("print", {"v"}, []() { return value(); });   // stdout side effect only; not part of the scene
		def_native("func", {"args", "body"}, [this]() -> value {
			value f; f.type = value::Func; f.fn = std::make_shared<func_t>();
			value args = lookup("args");
			for (const auto& a : args.items()) f.fn->argnames.push_back(a.var());
			f.fn->body = lookup("body");
			return f;
		});
		def_native("concat-all", {"vs"}, [this]() -> value {
			value_list out;
			value vs = lookup("vs");
			for (const auto& b : vs.items()) { const value_list& x = b.items(); out.insert(out.end(), x.begin(), x.end()); }
			return value::block(std::move(out));
		});
		def_native("append", {"block", "value"}, [this]() -> value {
			value_list v = lookup("block").items(); v.push_back(lookup("value")); return value::block(std::move(v));
		});
		def_native("collect-range", {"var", "range", "body"}, [this]() -> value {
			value rng = lookup("range"), body = lookup("body");
			int64_t a = rng.at(0).integer_(), b = rng.at(1).integer_();
			std::string id = lookup("var").ident();
			push();
			value_list out;
			for (int64_t k = a; k < b; ++k) { bind(id, value::integer(k)); out.push_back(eval(body)); }
			pop();
			return value::block(std::move(out));
		});
		def_native("block-format", {"format", "values"}, [this]() -> value {
			value_list fmt = lookup("format").items();
			value vals = reduce(lookup("values"));
			return format(fmt, vals.items());
		});
	}
	// resolve var / val / def / group; blocks stay unevaluated
	value eval1(const value& v) {
		switch (v.type) {
			case value::Group: {
				if (v.vs.empty()) throw parse_error("empty group");
				size_t i = 0;
				value rv = expr(v.vs, i);
				if (i + 1 < v.vs.size()) throw parse_error("a group holds exactly one expression");
				return rv;
			}
			case value::Var: case value::Val: return lookup(v.s);
			case value::Def: { value r = eval1(v.vs.at(0)); bind(v.s, r); return r; }
			case value::Native: return (*v.native)();
			default: return v;
		}
	}
	// `do`
	value eval(const value& v) {
		if (v.type != value::Block) return eval1(v);
		push();
		value rv;
		for (size_t i = 0; i < v.vs.size();) rv = expr(v.vs, i);
		pop();
		return rv;
	}
	// `reduce`
	value reduce(const value& v) {
		if (v.type != value::Block) return eval1(v);
		push();
		value_list out;
		for (size_t i = 0; i < v.vs.size();) out.push_back(expr(v.vs, i));
		pop();
		return value::block(std::move(out));
	}
};

// ---- scene builder (inc/scene.h:64-299) ---------------------------------------------------------------
struct v3 { float x, y, z; };
v3 operator-(v3 a, v3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
v3 operator+(v3 a, v3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
v3 operator*(float s, v3 a) { return {s * a.x, s * a.y, s * a.z}; }
v3 cross(v3 x, v3 y) { return {x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y}; }
float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
v3 normalize(v3 v) { float k = 1.f / std::sqrt(dot(v, v)); return {v.x * k, v.y * k, v.z * k}; }   // glm: v * inversesqrt(dot)
void put(float* d, v3 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }

// bk2v3(v): literal numbers; bk2v3(cx, v): `reduce` first (scene.h:18-26)
v3 block_to_v3(const value& v) { return {(float)v.at(0).num(), (float)v.at(1).num(), (float)v.at(2).num()}; }
v3 block_to_v3(eval_context& cx, const value& v) { value r = cx.reduce(v); return block_to_v3(r); }

struct mesh { std::vector<v3> pos, nrm; std::vector<float> uv; std::vector<std::array<uint32_t, 9>> faces; };

// the hand-rolled OBJ reader of scene.h:138-201: `v`, `vn`, `vt`, triangular `f a/b/c` records, 1-based
bool load_obj(const std::string& path, mesh& m) {
	std::ifstream inf(path);
	if (!inf) return false;
	std::string comm;
	while (inf >> comm) {
		if (comm == "v" || comm == "vn") { v3 p; inf >> p.x >> p.y >> p.z; (comm == "v" ? m.pos : m.nrm).push_back(p); }
		else if (comm == "vt") { float u, v; inf >> u >> v; m.uv.push_back(u); m.uv.push_back(v); }
		else if (comm == "f") {
			std::array<uint32_t, 9> f{};   // ip[3], it[3], in[3]
			for (int k = 0; k < 3; ++k) {
				uint32_t ip = 0, it = 0, in = 0;
				inf >> ip;
				if (inf.peek() == '/') {
					inf.ignore();
					if (inf.peek() != '/') inf >> it;
					if (inf.peek() == '/') { inf.ignore(); inf >> in; }
				}
				f[k] = ip - 1; f[3 + k] = it - 1; f[6 + k] = in - 1;
			}
			m.faces.push_back(f);
		}
	}
	return true;
}

struct builder {
	plu_scene& out;
	eval_context cx;
	// a named material is flattened when a surface first uses it (the reference's flattener discovers materials
	// by walking the surfaces, so unused definitions never reach the tables and the order is first-use order)
	struct pending { plu_material_desc m; bool has_tex; plu_texture_desc t; int32_t index; };
	std::map<std::string, pending> named_materials;
	std::string base_dir;
	explicit builder(plu_scene& s) : out(s) {}

	int32_t emit(pending& p) {
		if (p.index >= 0) return p.index;
		if (p.has_tex) { out.textures.push_back(p.t); p.m.texture = (int32_t)out.textures.size() - 1; }
		out.materials.push_back(p.m);
		return p.index = (int32_t)out.materials.size() - 1;
	}

	// make_color, scene.h:72-99: `[r g b]` or `texture [ checkerboard a b s | grid a b s l | img "file" ]`
	void make_color(const value_list& vs, size_t& i, pending& pm) {
		plu_material_desc& m = pm.m;
		if (i >= vs.size()) throw parse_error("material without a colour");
		if (vs[i].type == value::Var) {
			if (vs[i].s != "texture") throw parse_error("unknown colour source '" + vs[i].s + "'");
			i++;
			if (i >= vs.size()) throw parse_error("texture without a description block");
			const value_list& ts = vs[i].items(); i++;
			if (ts.empty()) throw parse_error("empty texture block");
			const std::string& kind = ts[0].var();
			plu_texture_desc t{};
			if (kind == "checkerboard" && ts.size() >= 4) {
				t.tag = PLU_TEX_CHECKERBOARD; put(t.c0, block_to_v3(cx, ts[1])); put(t.c1, block_to_v3(cx, ts[2])); t.scale = (float)cx.eval(ts[3]).num();
			} else if (kind == "grid" && ts.size() >= 5) {
				// grid_texture(fg, bg, scale, line) receives ts[1], ts[2] in that order (scene.h:84, textures.h:25)
				t.tag = PLU_TEX_GRID; put(t.c0, block_to_v3(cx, ts[1])); put(t.c1, block_to_v3(cx, ts[2]));
				t.scale = (float)cx.eval(ts[3]).num();
				value l = cx.eval(ts[4]); if (l.type != value::Float) throw parse_error("grid line size must be a Float"); t.line = (float)l.f;
			} else if (kind == "img" && ts.size() >= 2) {
				const std::string& file = ts[1].str();
				int32_t w = 0, h = 0;
				t.tag = PLU_TEX_IMAGE; t.img_offset = (int32_t)(out.image_texels.size() / 3);
				if (!load_bmp_texels(file, out.image_texels, w, h) && !load_bmp_texels(base_dir + file, out.image_texels, w, h))
					throw parse_error("cannot read BMP file '" + file + "'");
				t.img_w = w; t.img_h = h;
			} else throw parse_error("unknown texture '" + kind + "'");
			pm.has_tex = true; pm.t = t; m.texture = 0;
			m.color[0] = m.color[1] = m.color[2] = 0.f;   // props::color(tex) : col(0.f)
		} else if (vs[i].type == value::Block) {
			put(m.color, block_to_v3(cx, vs[i++]));
			m.texture = -1;
		} else throw parse_error("a colour is a block or `texture [...]`");
	}

	// make_material, scene.h:101-124
	pending make_material(const value& v) {
		const value_list& vs = v.items();
		if (vs.empty()) throw parse_error("empty material block");
		const std::string& kind = vs[0].var();
		pending pm{}; pm.index = -1; pm.has_tex = false;
		plu_material_desc& m = pm.m; m.texture = -1; m.light = -1;
		size_t i = 1;
		auto need = [&](size_t k) { if (i + k > vs.size()) throw parse_error("material '" + kind + "' is missing parameters"); };
		if (kind == "diffuse") { m.tag = PLU_MAT_DIFFUSE; make_color(vs, i, pm); }
		else if (kind == "perfect-reflection") {
			m.tag = PLU_MAT_PERFECT_REFLECTION; make_color(vs, i, pm); need(2);
			put(m.p0, block_to_v3(cx, vs[i++])); put(m.p1, block_to_v3(cx, vs[i++]));
		} else if (kind == "perfect-refraction") {
			m.tag = PLU_MAT_PERFECT_REFRACTION; make_color(vs, i, pm); need(2);
			m.p0[0] = (float)vs[i].num(); m.p0[1] = (float)vs[i + 1].num();   // (eta_t, eta_i), material.h:236
		} else if (kind == "glass") {
			m.tag = PLU_MAT_GLASS; make_color(vs, i, pm); need(1);
			m.p0[0] = (float)vs[i].num();
		} else throw parse_error("unknown material '" + kind + "'");
		return pm;
	}

	int32_t make_or_ref_material(const value& m) {   // scene.h:126-128
		if (m.type == value::Block) { pending p = make_material(m); return emit(p); }
		if (m.type == value::Id) {
			auto it = named_materials.find(m.s);
			if (it == named_materials.end()) throw parse_error("unknown material '" + m.s + "'");
			return emit(it->second);
		}
		throw parse_error("a surface must be followed by a material block or a 'name");
	}

	// make_basic_surface, scene.h:203-226: appends the surfaces (material -1 for now), returns how many
	size_t make_basic_surface(const value_list& vs, size_t& i) {
		const std::string& kind = vs[i].var();
		float d[12] = {0};
		if (kind == "sphere") {
			if (i + 2 > vs.size() - 1) throw parse_error("sphere needs a centre and a radius");
			put(d, block_to_v3(cx, vs[i + 1])); d[3] = (float)cx.eval(vs[i + 2]).num();
			i += 3;
			out.add_prim(PLU_PRIM_SPHERE, -1, d);
			return 1;
		}
		if (kind == "box") {
			if (i + 2 > vs.size() - 1) throw parse_error("box needs a centre and extents");
			v3 c = block_to_v3(cx, vs[i + 1]), e = block_to_v3(cx, vs[i + 2]);
			put(d, c - e); put(d + 4, c + e);   // box(c, e) : box_bounds(c - e, c + e), box.h:9
			i += 3;
			out.add_prim(PLU_PRIM_BOX, -1, d);
			return 1;
		}
		if (kind == "triangle-mesh") {
			if (i + 1 > vs.size() - 1) throw parse_error("triangle-mesh needs a file name");
			std::string path = vs[i + 1].str();
			i += 2;
			mesh m;
			if (!load_obj(path, m) && !load_obj(base_dir + path, m)) throw parse_error("cannot read OBJ file '" + path + "'");
			size_t n0 = out.prim_tag.size();
			out.prim_tri_uv.resize(n0 * 6, 0.f); out.prim_tri_normal.resize(n0 * 9, 0.f);
			for (const auto& f : m.faces) {
				for (int k = 0; k < 3; ++k) {
					if (f[k] >= m.pos.size()) throw parse_error("OBJ face refers to a missing vertex");
					put(d + 4 * k, m.pos[f[k]]);
				}
				out.add_prim(PLU_PRIM_TRIANGLE, -1, d);
				for (int k = 0; k < 3; ++k) {
					// a face without vt / vn reads uninitialised indices in the reference (scene.h:168-181); here: zeros
					float u = 0.f, v = 0.f; v3 nn{0.f, 0.f, 0.f};
					if ((size_t)f[3 + k] * 2 + 1 < m.uv.size()) { u = m.uv[(size_t)f[3 + k] * 2]; v = m.uv[(size_t)f[3 + k] * 2 + 1]; }
					if (f[6 + k] < m.nrm.size()) nn = m.nrm[f[6 + k]];
					out.prim_tri_uv.push_back(u); out.prim_tri_uv.push_back(v);
					out.prim_tri_normal.push_back(nn.x); out.prim_tri_normal.push_back(nn.y); out.prim_tri_normal.push_back(nn.z);
				}
			}
			return m.faces.size();
		}
		return 0;
	}

	// surface::area() as each class computes it (sphere.h:17 is the VOLUME formula; kept)
	float surface_area(size_t s) const {
		const float* f = &out.prim_data[s * 12];
		if (out.prim_tag[s] == PLU_PRIM_SPHERE) { float r = f[3]; return (4.f / 3.f) * 3.14159265358979323846264338327950288f * r * r * r; }
		if (out.prim_tag[s] == PLU_PRIM_BOX) { float x = f[4] - f[0], y = f[5] - f[1], z = f[6] - f[2]; return 2.f * (x * y + x * z + y * z); }
		auto dist = [&](int a, int b) { v3 d{f[4 * b] - f[4 * a], f[4 * b + 1] - f[4 * a + 1], f[4 * b + 2] - f[4 * a + 2]}; return std::sqrt(dot(d, d)); };
		float a = dist(0, 1), b = dist(1, 2), c = dist(0, 2);   // triangle.h:25-31 (Heron)
		float p = (a + b + c) * 0.5f;
		return std::sqrt(p * (p - a) * (p - b) * (p - c));
	}

	void run(const value& top, std::list<std::string>& args) {
		// resolution + "/res WxH" (scene.h:230-238)
		value res = top.named("resolution");
		uint32_t W = 1280, H = 960;
		if (!res.is_null()) { W = (uint32_t)res.at(0).integer_(); H = (uint32_t)res.at(1).integer_(); }
		for (auto it = args.begin(); it != args.end(); ++it) if (*it == "/res") {
			auto nx = std::next(it);
			if (nx == args.end()) throw parse_error("/res needs WxH");
			std::string r = *nx; size_t x = r.find('x');
			W = (uint32_t)std::atoi(r.substr(0, x).c_str()); H = x == std::string::npos ? 0u : (uint32_t)std::atoi(r.substr(x + 1).c_str());
			break;
		}
		// camera (scene.h:240-248, camera.h:17-23): `up:` is never read; both axes scaled by 1.5, w = 2.5
		value cam = top.named("camera");
		if (cam.is_null()) throw parse_error("scene has no camera block");
		float lens_r = 0.f, focal = 0.f;
		if (cam.has_named("lens")) { value l = cam.named("lens"); lens_r = (float)l.named("radius").num(); focal = (float)l.named("focal-distance").num(); }
		v3 pos = block_to_v3(cam.named("position")), target = block_to_v3(cam.named("target"));
		v3 look = normalize(target - pos);
		v3 right = 1.5f * normalize(cross(look, v3{0.f, -1.f, 0.f}));
		v3 up = 1.5f * normalize(cross(look, right));
		put(out.cam.pos, pos); put(out.cam.look, look); put(out.cam.up, up); put(out.cam.right, right);
		out.cam.inv_image_size[0] = 1.f / (float)W; out.cam.inv_image_size[1] = 1.f / (float)H;
		out.cam.w = 2.5f; out.cam.lens_radius = lens_r; out.cam.focal_distance = focal;
		// antialiasing-samples + "/smp N" (scene.h:250-256)
		int64_t samples = top.named("antialiasing-samples").integer_();
		for (auto it = args.begin(); it != args.end(); ++it) if (*it == "/smp") {
			auto nx = std::next(it);
			if (nx == args.end()) throw parse_error("/smp needs a number");
			samples = std::atoll(nx->c_str());
			break;
		}
		out.width = (int32_t)W; out.height = (int32_t)H; out.n_strata = (int32_t)samples;

		// materials: a block of definitions (scene.h:261-266)
		value mats = top.named("materials");
		for (const auto& mv : mats.items()) {
			if (mv.type != value::Def) throw parse_error("the materials block takes `name: [ ... ]` definitions only");
			named_materials[mv.s] = make_material(mv.vs.at(0));
		}
		// objects (scene.h:268-297): resolved once (a group is evaluated, a block is taken as is), then a flat walk
		value objs = cx.eval1(top.named("objects"));
		const value_list& ov = objs.items();
		for (size_t i = 0; i < ov.size();) {
			size_t first = out.prim_tag.size();
			size_t made = make_basic_surface(ov, i);
			if (made == 0) {
				const std::string& kind = ov[i].var();
				if (kind == "point-light") {
					if (i + 2 > ov.size() - 1) throw parse_error("point-light needs a position and an intensity");
					plu_light_desc l{}; l.tag = PLU_LIGHT_POINT; l.prim = -1;
					put(l.a, block_to_v3(cx, ov[i + 1])); put(l.b, block_to_v3(cx, ov[i + 2]));
					out.lights.push_back(l);
					i += 3;
				} else if (kind == "diffuse-area-light") {
					if (i + 2 > ov.size() - 1) throw parse_error("diffuse-area-light needs a surface block and a radiance");
					value sb = cx.eval1(ov[i + 1]);
					size_t j = 0, s0 = out.prim_tag.size();
					if (sb.items().empty() || make_basic_surface(sb.items(), j) != 1) throw parse_error("diffuse-area-light takes exactly one surface");
					plu_light_desc l{}; l.tag = PLU_LIGHT_DIFFUSE_AREA; l.prim = (int32_t)s0;
					put(l.b, block_to_v3(cx, ov[i + 2]));
					l.area = surface_area(s0);
					plu_material_desc m{}; m.tag = PLU_MAT_EMISSION; m.texture = -1; m.light = (int32_t)out.lights.size();
					out.materials.push_back(m);
					out.prim_material[s0] = (int32_t)out.materials.size() - 1;
					out.lights.push_back(l);
					i += 3;
				} else throw parse_error("unknown object '" + kind + "'");
			} else {
				if (i >= ov.size()) throw parse_error("surface without a material");
				int32_t m = make_or_ref_material(ov[i]);
				i++;
				for (size_t s = first; s < first + made; ++s) out.prim_material[s] = m;
			}
		}
		if (!out.prim_tri_uv.empty()) { out.prim_tri_uv.resize(out.prim_tag.size() * 6, 0.f); out.prim_tri_normal.resize(out.prim_tag.size() * 9, 0.f); }
	}
};

}  // namespace

extern "C" int plu_scene_load_urn(const char* path, const char* const* args, int n_args, plu_scene** out, char* err, int err_len) {
	if (!path || !out || n_args < 0 || (n_args > 0 && !args)) return PLU_E_ARG;
	*out = nullptr;
	std::ifstream in(path);
	if (!in) { pluhost::set_err(err, err_len, std::string("cannot open ") + path); return PLU_E_IO; }
	try {
		token_stream ts(in);
		value top = parse(ts);
		if (top.type != value::Block) throw parse_error("a scene file is one top-level [ block ]");
		std::unique_ptr<plu_scene> s(new plu_scene);
		builder b(*s);
		std::string p(path); size_t sl = p.find_last_of('/');
		b.base_dir = sl == std::string::npos ? "" : p.substr(0, sl + 1);
		std::list<std::string> a;
		for (int i = 0; i < n_args; ++i) a.push_back(args[i] ? args[i] : "");
		b.run(top, a);
		s->seal();
		{   // tie ranks of the reference tree (bvh_tree.cpp:7-36), so the scene is complete without a live bvh_tree
			std::vector<int32_t> rank;
			pluhost::reference_leaf_order(&s->desc, rank);
			s->prim_tie_rank = rank;
			s->seal();
		}
		std::string v = pluhost::validate_desc(&s->desc);
		if (!v.empty()) throw parse_error(v);
		*out = s.release();
		return PLU_OK;
	} catch (const std::exception& e) {
		pluhost::set_err(err, err_len, std::string(path) + ": " + e.what());
		return PLU_E_PARSE;
	}
}
