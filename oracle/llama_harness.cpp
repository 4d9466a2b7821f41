// oracle/llama_harness.cpp — TEST INFRASTRUCTURE ONLY (not product code).
//
// Greedy-decode token-stream parity through the REAL llama.cpp graph and loaders:
//   gen : writes a random-init Llama-shaped GGUF with the reference's own gguf API (ggml/include/gguf.h) - every matmul weight
//         quantized by ggml_quantize_chunk (ggml.c:7442), tokenizer.ggml.model = "no_vocab" (LLAMA_VOCAB_TYPE_NONE,
//         src/llama-vocab.cpp:1692) - as one file, as `n` split files (the layout tools/gguf-split writes, gguf-split.cpp:225-380)
//         and the <prefix>.tensors.txt list the split-future loader wants.
//   run : loads it with llama_model_load_from_file / llama_model_load_from_buffer / llama_model_load_from_split_futures
//         (src/llama.cpp:313, 328, 361), runs llama_decode on a prompt and then token by token with argmax sampling, and prints
//         the token stream with the top-2 logit margin of every step as one JSON line.
// Backends are loaded explicitly with ggml_backend_load (the reference's registry): a ggml-cpu variant always, and
// libggml-b200.so when the run is to be offloaded (-ngl 99). Built against the reference's headers where they lie and linked to
// oracle/_ref/libllama.so; nothing here is compiled into the product.

#include "llama.h"
#include "llama-cpp.h"
#include "ggml.h"
#include "ggml-backend.h"
#include "gguf.h"
#include "uint8-buff-stream.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <thread>
#include <vector>

namespace {

struct model_shape {
    int n_layer = 4, n_embd = 4096, n_ff = 14336, n_head = 32, n_head_kv = 8, n_vocab = 8192, n_ctx = 1024;
};

uint64_t splitmix(uint64_t & s) {
    uint64_t z = (s += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

// row r of tensor t: U(-a, a), the same values whatever the thread count
void fill_row(float * dst, int64_t n, uint64_t seed, int tensor, int64_t row, float a) {
    uint64_t s = seed * 0x100000001b3ull + (uint64_t) tensor * 0x9e3779b1ull + (uint64_t) row * 0x85ebca77c2b2ae63ull;
    for (int64_t i = 0; i < n; i++) {
        const uint64_t r = splitmix(s);
        dst[i] = ((float) (r >> 40) * (1.0f / 8388608.0f) - 1.0f) * a;  // 24 bits -> [-1, 1)
    }
}

int n_workers() {
    const unsigned hc = std::thread::hardware_concurrency();
    return (int) std::max(1u, std::min(hc ? hc : 4u, 64u));
}

ggml_tensor * make_tensor(ggml_context * ctx, const std::string & name, ggml_type type, int64_t ne0, int64_t ne1, uint64_t seed, int index, float amp,
                          bool ones) {
    ggml_tensor * t = ne1 > 1 ? ggml_new_tensor_2d(ctx, type, ne0, ne1) : ggml_new_tensor_1d(ctx, type, ne0);
    ggml_set_name(t, name.c_str());
    if (ones) {
        for (int64_t i = 0; i < ne0; i++) ((float *) t->data)[i] = 1.0f;
        return t;
    }
    const int nw = n_workers();
    std::vector<std::thread> th;
    const size_t row_size = ggml_row_size(type, ne0);
    for (int w = 0; w < nw; w++) {
        th.emplace_back([=]() {
            const int64_t r0 = ne1 * w / nw, r1 = ne1 * (w + 1) / nw;
            std::vector<float> buf((size_t) ne0);
            for (int64_t r = r0; r < r1; r++) {
                fill_row(buf.data(), ne0, seed, index, r, amp);
                if (type == GGML_TYPE_F32) memcpy((char *) t->data + r * row_size, buf.data(), row_size);
                else ggml_quantize_chunk(type, buf.data(), (char *) t->data + r * row_size, 0, 1, ne0, nullptr);
            }
        });
    }
    for (auto & x : th) x.join();
    return t;
}

void set_model_kv(gguf_context * g, const model_shape & m, ggml_type wtype) {
    gguf_set_val_str(g, "general.architecture", "llama");
    gguf_set_val_str(g, "general.name", "b200-parity-random-init");
    gguf_set_val_u32(g, "llama.context_length", (uint32_t) m.n_ctx);
    gguf_set_val_u32(g, "llama.embedding_length", (uint32_t) m.n_embd);
    gguf_set_val_u32(g, "llama.block_count", (uint32_t) m.n_layer);
    gguf_set_val_u32(g, "llama.feed_forward_length", (uint32_t) m.n_ff);
    gguf_set_val_u32(g, "llama.attention.head_count", (uint32_t) m.n_head);
    gguf_set_val_u32(g, "llama.attention.head_count_kv", (uint32_t) m.n_head_kv);
    gguf_set_val_f32(g, "llama.attention.layer_norm_rms_epsilon", 1e-5f);
    gguf_set_val_u32(g, "llama.rope.dimension_count", (uint32_t) (m.n_embd / m.n_head));
    gguf_set_val_f32(g, "llama.rope.freq_base", 500000.0f);
    gguf_set_val_u32(g, "llama.vocab_size", (uint32_t) m.n_vocab);
    gguf_set_val_str(g, "tokenizer.ggml.model", "no_vocab");
    gguf_set_val_u32(g, "general.quantization_version", 2);
    (void) wtype;
}

bool write_gguf(gguf_context * g, const std::string & path) {
    if (!gguf_write_to_file(g, path.c_str(), false)) {
        fprintf(stderr, "llama_harness: cannot write %s\n", path.c_str());
        return false;
    }
    return true;
}

int cmd_gen(int argc, char ** argv) {
    model_shape m;
    std::string out = "/tmp/b200_parity", tname = "q4_K";
    uint64_t seed = 1;
    int n_split = 3;
    for (int i = 0; i < argc; i++) {
        const std::string a = argv[i];
        auto val = [&]() { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--out") out = val();
        else if (a == "--type") tname = val();
        else if (a == "--layers") m.n_layer = atoi(val());
        else if (a == "--embd") m.n_embd = atoi(val());
        else if (a == "--ff") m.n_ff = atoi(val());
        else if (a == "--heads") m.n_head = atoi(val());
        else if (a == "--kv-heads") m.n_head_kv = atoi(val());
        else if (a == "--vocab") m.n_vocab = atoi(val());
        else if (a == "--ctx") m.n_ctx = atoi(val());
        else if (a == "--seed") seed = strtoull(val(), nullptr, 10);
        else if (a == "--splits") n_split = atoi(val());
    }
    ggml_type wtype = GGML_TYPE_COUNT;
    for (int t = 0; t < GGML_TYPE_COUNT; t++) if (tname == ggml_type_name((ggml_type) t)) wtype = (ggml_type) t;
    if (wtype == GGML_TYPE_COUNT) { fprintf(stderr, "unknown type %s\n", tname.c_str()); return 2; }
    const int64_t E = m.n_embd, F = m.n_ff, KV = (int64_t) m.n_embd / m.n_head * m.n_head_kv, V = m.n_vocab;

    struct spec { std::string name; ggml_type type; int64_t ne0, ne1; bool ones; };
    std::vector<spec> specs;
    specs.push_back({ "token_embd.weight", GGML_TYPE_F32, E, V, false });
    for (int l = 0; l < m.n_layer; l++) {
        const std::string p = "blk." + std::to_string(l) + ".";
        specs.push_back({ p + "attn_norm.weight", GGML_TYPE_F32, E, 1, true });
        specs.push_back({ p + "attn_q.weight", wtype, E, E, false });
        specs.push_back({ p + "attn_k.weight", wtype, E, KV, false });
        specs.push_back({ p + "attn_v.weight", wtype, E, KV, false });
        specs.push_back({ p + "attn_output.weight", wtype, E, E, false });
        specs.push_back({ p + "ffn_norm.weight", GGML_TYPE_F32, E, 1, true });
        specs.push_back({ p + "ffn_gate.weight", wtype, E, F, false });
        specs.push_back({ p + "ffn_up.weight", wtype, E, F, false });
        specs.push_back({ p + "ffn_down.weight", wtype, F, E, false });
    }
    specs.push_back({ "output_norm.weight", GGML_TYPE_F32, E, 1, true });
    specs.push_back({ "output.weight", wtype, E, V, false });

    size_t bytes = 0;
    for (auto & s : specs) bytes += ggml_row_size(s.type, s.ne0) * (size_t) s.ne1 + 64;
    ggml_init_params ip = { bytes + ggml_tensor_overhead() * (specs.size() + 8), nullptr, false };
    ggml_context * ctx = ggml_init(ip);
    if (!ctx) { fprintf(stderr, "ggml_init failed (%zu bytes)\n", bytes); return 1; }
    std::vector<ggml_tensor *> tensors;
    for (size_t i = 0; i < specs.size(); i++) {
        const spec & s = specs[i];
        // variance-preserving range for the matrices, so that the residual stream and the logits stay O(1)
        const float amp = s.name == "token_embd.weight" ? 1.0f : sqrtf(3.0f / (float) s.ne0);
        tensors.push_back(make_tensor(ctx, s.name, s.type, s.ne0, s.ne1, seed, (int) i, amp, s.ones));
    }

    // one file
    {
        gguf_context * g = gguf_init_empty();
        set_model_kv(g, m, wtype);
        for (ggml_tensor * t : tensors) gguf_add_tensor(g, t);
        const bool ok = write_gguf(g, out + ".gguf");
        gguf_free(g);
        if (!ok) return 1;
    }
    // n_split files the way tools/gguf-split lays them out: all model KV in split 0, split.no / split.count / split.tensors.count in each
    if (n_split > 1) {
        std::ofstream list(out + ".tensors.txt");
        for (ggml_tensor * t : tensors) list << t->name << "\n";
        list.close();
        const int n = (int) tensors.size();
        for (int s = 0; s < n_split; s++) {
            gguf_context * g = gguf_init_empty();
            if (s == 0) set_model_kv(g, m, wtype);
            gguf_set_val_u16(g, "split.no", (uint16_t) s);
            gguf_set_val_u16(g, "split.count", (uint16_t) n_split);
            gguf_set_val_i32(g, "split.tensors.count", n);
            for (int i = n * s / n_split; i < n * (s + 1) / n_split; i++) gguf_add_tensor(g, tensors[i]);
            char path[1024];
            llama_split_path(path, sizeof(path), out.c_str(), s, n_split);
            const bool ok = write_gguf(g, path);
            gguf_free(g);
            if (!ok) return 1;
        }
    }
    printf("{\"out\": \"%s\", \"type\": \"%s\", \"tensors\": %zu, \"bytes\": %zu, \"splits\": %d}\n", out.c_str(), tname.c_str(), tensors.size(), bytes, n_split);
    ggml_free(ctx);
    return 0;
}

std::vector<uint8_t> slurp(const std::string & path) {
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) { fprintf(stderr, "llama_harness: cannot read %s\n", path.c_str()); exit(1); }
    const size_t n = (size_t) f.tellg();
    f.seekg(0);
    std::vector<uint8_t> b(n);
    f.read((char *) b.data(), (std::streamsize) n);
    return b;
}

int g_graph_splits = -1;   // "graph splits = N" of llama_context's reserve pass

int cmd_run(int argc, char ** argv) {
    std::string prefix = "/tmp/b200_parity", loader = "file";
    std::vector<std::string> backends;
    int ngl = 0, n_prompt = 64, n_gen = 256, n_threads = 0, n_split = 3, n_batch = 512, n_ubatch = 512;
    uint64_t seed = 5;
    bool quiet = true, mmap = true, offload_kqv = false, flash_attn = false;
    int n_warm = 0;
    std::string force_path;   // teacher forcing: a file of token ids; token i of it is fed after step i instead of this run's own argmax
    for (int i = 0; i < argc; i++) {
        const std::string a = argv[i];
        auto val = [&]() { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--model") prefix = val();
        else if (a == "--loader") loader = val();
        else if (a == "--backend") backends.push_back(val());
        else if (a == "--ngl") ngl = atoi(val());
        else if (a == "--prompt") n_prompt = atoi(val());
        else if (a == "--gen") n_gen = atoi(val());
        else if (a == "--threads") n_threads = atoi(val());
        else if (a == "--splits") n_split = atoi(val());
        else if (a == "--batch") n_batch = atoi(val());
        else if (a == "--ubatch") n_ubatch = atoi(val());
        else if (a == "--seed") seed = strtoull(val(), nullptr, 10);
        else if (a == "--verbose") quiet = false;
        else if (a == "--no-mmap") mmap = false;
        else if (a == "--force") force_path = val();
        else if (a == "--warmup") n_warm = atoi(val());   // untimed: the prompt + this many tokens first, then the cache is cleared (llama-bench's warmup run)
        else if (a == "--kqv") offload_kqv = true;     // KV cache and attention on the device of the layer
        else if (a == "--fa") flash_attn = true;       // GGML_OP_FLASH_ATTN_EXT instead of MUL_MAT / SOFT_MAX / MUL_MAT
    }
    llama_log_set([](ggml_log_level level, const char * text, void * ud) {
        const char * gs = strstr(text, "graph splits = ");   // llama_context reports the scheduler's split count of its reserve graphs
        if (gs) g_graph_splits = atoi(gs + 15);
        if (level >= GGML_LOG_LEVEL_ERROR || !*(bool *) ud) fputs(text, stderr);
    }, &quiet);
    for (const std::string & b : backends) {
        if (!ggml_backend_load(b.c_str())) { fprintf(stderr, "llama_harness: cannot load backend %s\n", b.c_str()); return 1; }
    }
    llama_backend_init();
    if (n_threads <= 0) n_threads = n_workers();

    llama_model_params mp = llama_model_default_params();
    mp.n_gpu_layers = ngl;
    mp.use_mmap     = mmap;
    const auto t_load0 = std::chrono::steady_clock::now();
    llama_model * model = nullptr;
    if (loader == "file") {
        model = llama_model_load_from_file((prefix + ".gguf").c_str(), mp);
    } else if (loader == "buffer") {
        model = llama_model_load_from_buffer(slurp(prefix + ".gguf"), mp);
    } else if (loader == "futures") {
        std::vector<std::string> paths;
        for (int s = 0; s < n_split; s++) {
            char path[1024];
            llama_split_path(path, sizeof(path), prefix.c_str(), s, n_split);
            paths.push_back(path);
        }
        const std::string list_path = prefix + ".tensors.txt";
        std::vector<const char *> cpaths;
        for (auto & p : paths) cpaths.push_back(p.c_str());
        const char * lctx = "b200-parity";
        std::thread fulfill([&]() {  // the files arrive from another thread while the loader is already waiting (tests/test-model-load-memory-split.cpp)
            if (!llama_model_load_fulfill_split_future(list_path.c_str(), lctx, std::make_unique<Uint8BufferStreamBuf>(slurp(list_path)))) exit(3);
            for (auto & p : paths) {
                if (!llama_model_load_fulfill_split_future(p.c_str(), lctx, std::make_unique<Uint8BufferStreamBuf>(slurp(p)))) exit(3);
            }
        });
        model = llama_model_load_from_split_futures(cpaths.data(), cpaths.size(), lctx, list_path.c_str(), mp);
        fulfill.join();
    } else {
        fprintf(stderr, "unknown loader %s\n", loader.c_str());
        return 2;
    }
    if (!model) { fprintf(stderr, "llama_harness: model load failed (%s)\n", loader.c_str()); return 1; }
    const double load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_load0).count();

    llama_context_params cp = llama_context_default_params();
    cp.n_ctx           = (uint32_t) (n_prompt + n_gen + 8);
    cp.n_batch         = (uint32_t) n_batch;
    cp.n_ubatch        = (uint32_t) n_ubatch;
    cp.n_threads       = n_threads;
    cp.n_threads_batch = n_threads;
    cp.offload_kqv     = offload_kqv;  // default false: the KV cache and the attention ops stay on the CPU backend (the backend under test owns the matmuls only)
    cp.no_perf         = true;
    cp.flash_attn_type = flash_attn ? LLAMA_FLASH_ATTN_TYPE_ENABLED : LLAMA_FLASH_ATTN_TYPE_DISABLED;  // never AUTO: one attention arithmetic whatever the devices
    llama_context * ctx = llama_init_from_model(model, cp);
    if (!ctx) { fprintf(stderr, "llama_harness: context creation failed\n"); return 1; }
    const int n_vocab = llama_vocab_n_tokens(llama_model_get_vocab(model));

    std::vector<llama_token> toks;
    uint64_t s = seed;
    for (int i = 0; i < n_prompt; i++) toks.push_back((llama_token) (splitmix(s) % (uint64_t) n_vocab));
    std::vector<llama_token> forced;
    if (!force_path.empty()) {
        std::ifstream f(force_path);
        long v;
        while (f >> v) forced.push_back((llama_token) v);
        if ((int) forced.size() < n_gen) { fprintf(stderr, "llama_harness: %s holds %zu tokens, need %d\n", force_path.c_str(), forced.size(), n_gen); return 2; }
    }
    std::vector<llama_token> out;
    std::vector<float> margins, deficits;   // deficit: this run's top logit minus its logit of the forced token (0 where they agree)
    auto pick = [&]() {
        const float * lg = llama_get_logits_ith(ctx, -1);
        int best = 0;
        float b1 = -INFINITY, b2 = -INFINITY;
        for (int v = 0; v < n_vocab; v++) {
            if (lg[v] > b1) { b2 = b1; b1 = lg[v]; best = v; }
            else if (lg[v] > b2) b2 = lg[v];
        }
        margins.push_back(b1 - b2);
        if (!forced.empty()) deficits.push_back(b1 - lg[forced[margins.size() - 1]]);
        return (llama_token) best;
    };
    auto decode_prompt = [&]() {   // in pieces of n_batch, as llama-bench feeds a prompt (llama_context::decode asserts n_tokens <= n_batch)
        for (size_t i0 = 0; i0 < toks.size(); i0 += (size_t) n_batch) {
            const int32_t n = (int32_t) std::min(toks.size() - i0, (size_t) n_batch);
            if (llama_decode(ctx, llama_batch_get_one(toks.data() + i0, n)) != 0) return false;
        }
        return true;
    };
    if (n_warm > 0) {
        if (!decode_prompt()) { fprintf(stderr, "llama_decode(warmup prompt) failed\n"); return 1; }
        llama_token w = toks[0];
        for (int i = 0; i < n_warm; i++) {
            if (llama_decode(ctx, llama_batch_get_one(&w, 1)) != 0) { fprintf(stderr, "llama_decode(warmup token) failed\n"); return 1; }
        }
        llama_synchronize(ctx);
        llama_memory_clear(llama_get_memory(ctx), true);
    }
    const auto t0 = std::chrono::steady_clock::now();
    if (!decode_prompt()) { fprintf(stderr, "llama_decode(prompt) failed\n"); return 1; }
    llama_synchronize(ctx);   // device backends return from llama_decode before the work is done
    const double prompt_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    llama_token cur = pick();
    out.push_back(cur);
    if (!forced.empty()) cur = forced[0];
    const auto t1 = std::chrono::steady_clock::now();
    for (int i = 1; i < n_gen; i++) {
        if (llama_decode(ctx, llama_batch_get_one(&cur, 1)) != 0) { fprintf(stderr, "llama_decode(token %d) failed\n", i); return 1; }
        cur = pick();
        out.push_back(cur);
        if (!forced.empty()) cur = forced[i];
    }
    const double gen_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count();

    printf("{\"loader\": \"%s\", \"ngl\": %d, \"n_prompt\": %d, \"load_s\": %.3f, \"prompt_s\": %.6f, \"gen_s\": %.6f, \"graph_splits\": %d, \"tokens\": [", loader.c_str(), ngl,
           n_prompt, load_s, prompt_s, gen_s, g_graph_splits);
    for (size_t i = 0; i < out.size(); i++) printf("%s%d", i ? ", " : "", out[i]);
    printf("], \"margins\": [");
    for (size_t i = 0; i < margins.size(); i++) printf("%s%.6g", i ? ", " : "", margins[i]);
    printf("], \"deficits\": [");
    for (size_t i = 0; i < deficits.size(); i++) printf("%s%.6g", i ? ", " : "", deficits[i]);
    printf("]}\n");
    fflush(stdout);
    llama_free(ctx);
    llama_model_free(model);
    llama_backend_free();
    return 0;
}

} // namespace

int main(int argc, char ** argv) {
    if (argc >= 2 && !strcmp(argv[1], "gen")) return cmd_gen(argc - 2, argv + 2);
    if (argc >= 2 && !strcmp(argv[1], "run")) return cmd_run(argc - 2, argv + 2);
    fprintf(stderr, "usage: llama_harness gen --type q4_K --out PREFIX [--layers 4 --embd 4096 --ff 14336 --heads 32 --kv-heads 8 --vocab 8192 --splits 3 --seed 1]\n"
                    "       llama_harness run --model PREFIX --loader file|buffer|futures --backend LIB.so [--backend LIB2.so] [--ngl 99] [--prompt 64 --gen 256]\n");
    return 2;
}
