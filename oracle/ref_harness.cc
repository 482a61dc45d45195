// oracle/ref_harness.cc -- TEST INFRASTRUCTURE ONLY.
//
// A small main of ours around the UNMODIFIED reference host code (linked from
// /root/reference by oracle/Makefile). It runs one of the reference driver's seven
// operator chains (token_test.cc:1003,1188,1253,1319,1393,1458,1518 -- the reference's own
// test_spmm_* functions, compiled with -Dmain=ref_token_test_main) on a .mtx file and
// dumps EVERY item of the resulting meta_data_set as text, one value per line, via the
// reference's own writer (meta_data_item::output_2_file, metadata_set.cc:193).
//
// The reference's execute_binary (executor.cc:6) would fork a compiler on the generated
// program, which cannot compile in this snapshot; this file
// supplies execute_binary instead (executor.cc is left out of the link) and uses the call
// as the hook "the whole chain has run" to dump the metadata set.
//
// usage: ref_harness <family> <matrix.mtx> <N> <sparse_coarsen> <coarsen> <dump_dir>
//        ref_harness "ops=<name>[,<arg>...];<name>..." <matrix.mtx> <N> 0 0 <dump_dir>
//   the second form runs an arbitrary sequence of the reference's own operators (operator.hpp, the
//   code_generator constructors; arguments in declaration order, bools as 0/1) the way the driver chains do
//   (operator_executer::add_and_run), prints each operator's records and dumps the set at the end -- for the
//   operators none of the seven chains reaches (merge path, balanced intervals, column / nnz direction ...).
//   CWD must hold global_config.json, arm_code/ and data_source/ (tests/golden/gen_golden.py
//   prepares such a scratch dir).
#include "metadata_set.hpp"
#include "config.hpp"
#include "executor.hpp"
#include "operator.hpp"
#include "operator_executer.hpp"
#include "code_generator.hpp"
#include <iostream>
#include <sstream>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include <sys/stat.h>

using namespace std;

float test_spmm_thread_total(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_warp_total(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_warp_bitmap(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_thread_bit_map(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_warp_segment(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_block_total(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_block_bitmap(shared_ptr<meta_data_set>, string, int, int);

static shared_ptr<meta_data_set> g_set;
static string g_dump_dir;

// every key the seven chains can leave behind; meta_data_set has no public iterator, so probe
static const char *k_names[] = {
    "nz_row_indices", "nz_col_indices", "nz_vals", "origin_row_num", "origin_col_num", "origin_nnz_num",
    "begin_row_index", "end_row_index", "begin_col_index", "end_col_index", "original_nz_row_indices",
    "first_row_indices", "first_nz_indices", "first_row_indices_without_ending",
    "first_row_indices_relative_to_BMTB", "first_nz_indices_relative_to_BMTB",
    "first_row_indices_relative_to_BMW", "first_nz_indices_relative_to_BMW",
    "first_BMW_indices", "first_BMT_indices", "BMT_size_of_each_blk", "BMW_size_of_each_blk", "BMTB_size", "BMTB_size_of_each_blk",
    "nz_row_indices_after_interlance_storage", "nz_col_indices_after_interlance_storage",
    "nz_vals_after_interlance_storage", "thread_bit_map", "bit_map_of_thread", "segment_ptr", "segment_offset",
    "segment_empty_flag", "segment_empty_row_indices", "first_row_indices_without_ending_relative_to_BMTB",
    "first_row_indices_without_ending_relative_to_BMW"};
static const POS_TYPE k_pos[] = {GLOBAL_META, TBLOCK_META, WARP_META, THREAD_META};

static void dump_all()
{
    mkdir(g_dump_dir.c_str(), 0755);
    FILE *idx = fopen((g_dump_dir + "/INDEX").c_str(), "w");
    for (POS_TYPE pos : k_pos)
        for (const char *name : k_names)
            for (int id = -1; id <= 48; id++) // sub-matrix ids the division operators create
            {
                if (!g_set->is_exist(pos, name, id))
                    continue;
                string key = get_metadata_item_name(pos, name, id);
                shared_ptr<meta_data_item> item = g_set->get_element(pos, name, id);
                item->output_2_file(g_dump_dir + "/" + key);
                shared_ptr<universal_array> arr = item->get_metadata_arr();
                // key, length, in-memory dtype, compressed dtype (code_source_data.cc:383)
                fprintf(idx, "%s %lu %s %s\n", key.c_str(), (unsigned long)arr->get_len(),
                        convert_data_type_to_string(arr->get_data_type()).c_str(),
                        convert_data_type_to_string(arr->get_compress_data_type()).c_str());
            }
    fclose(idx);
}

bool execute_binary(string execute_path, string filename, float &total_execute_time, float &gflops)
{
    // remember where the reference wrote its own hand-off (registered arrays + kernel_file.cpp)
    FILE *f = fopen((g_dump_dir + ".handoff").c_str(), "w");
    fprintf(f, "%s\n", execute_path.c_str());
    fclose(f);
    dump_all();
    total_execute_time = 0;
    gflops = 0;
    return true;
}

bool compile_spmv_code(string) { return true; }

// one operator of the reference by name; a = the constructor's arguments between code_generator_ptr and operator_history
static shared_ptr<basic_operator> ref_operator(const string &op, shared_ptr<code_generator> cg, shared_ptr<operator_context> h, const vector<long> &a)
{
    auto A = [&](size_t i) -> long { if (i >= a.size()) { fprintf(stderr, "%s: missing argument %zu\n", op.c_str(), i); exit(2); } return a[i]; };
#define OP0(T) if (op == #T) return shared_ptr<basic_operator>(new T(cg, h))
#define OP1(T) if (op == #T) return shared_ptr<basic_operator>(new T(cg, A(0), h))
#define OP2(T) if (op == #T) return shared_ptr<basic_operator>(new T(cg, A(0), A(1), h))
#define OP3(T) if (op == #T) return shared_ptr<basic_operator>(new T(cg, A(0), A(1), A(2), h))
#define OP4(T) if (op == #T) return shared_ptr<basic_operator>(new T(cg, A(0), A(1), A(2), A(3), h))
#define OP5(T) if (op == #T) return shared_ptr<basic_operator>(new T(cg, A(0), A(1), A(2), A(3), A(4), h))
    OP0(sort_operator);
    OP1(fixed_interval_row_matrix_div_operator);
    OP3(row_nz_matrix_div_operator);
    OP2(fixed_interval_row_direction_tblock_blocking_operator);
    OP4(fixed_interval_row_direction_warp_blocking_operator);
    OP0(empty_row_pad_operator);
    OP5(fixed_interval_col_direction_thread_blocking_operator);
    OP0(interlance_storage_operator);
    if (op == "fixed_interval_row_direction_thread_blocking_operator")
        return shared_ptr<basic_operator>(new fixed_interval_row_direction_thread_blocking_operator(cg, A(0), A(1), A(2), A(3), A(4), A(5), A(6), h));
    OP5(fixed_interval_col_direction_warp_blocking_operator);
    OP3(fixed_interval_col_direction_tblock_blocking_operator);
    OP3(balanced_interval_row_direction_thread_blocking_operator);
    OP3(balanced_interval_row_direction_warp_blocking_operator);
    OP1(balanced_interval_row_direction_tblock_blocking_operator);
    OP4(fixed_interval_nnz_direction_thread_blocking_operator);
    OP4(fixed_interval_nnz_direction_warp_blocking_operator);
    OP2(fixed_interval_nnz_direction_tblock_blocking_operator);
    OP1(merge_path_tblock_operator);
    OP1(merge_path_warp_operator);
    OP1(merge_path_thread_operator);
    fprintf(stderr, "ops mode: unknown operator %s\n", op.c_str());
    exit(2);
}

static int run_ops(const string &spec)
{
    shared_ptr<code_generator> cg(new code_generator(g_set, 0));
    shared_ptr<operator_executer> ex(new operator_executer());
    stringstream ss(spec);
    string item;
    while (getline(ss, item, ';'))
    {
        if (item.empty()) continue;
        stringstream is(item);
        string name, tok;
        getline(is, name, ',');
        vector<long> a;
        while (getline(is, tok, ',')) a.push_back(atol(tok.c_str()));
        shared_ptr<basic_operator> o = ref_operator(name, cg, ex->get_operator_context(), a);
        // the verdict of the operator's own rule on the history so far (the search uses it to prune chains)
        cout << "valid::" << name << "::" << (o->is_valid_according_to_operator(ex->get_operator_context()) ? 1 : 0) << endl;
        ex->add_and_run(o);
        cout << o->convert_to_string() << endl;
        vector<shared_ptr<transform_step_record_item>> rec = o->get_data_transform_sequence();
        for (size_t i = 0; i < rec.size(); i++) cout << rec[i]->convert_to_string() << endl;
    }
    dump_all();
    return 0;
}

int main(int argc, char **argv)
{
    if (argc != 7)
    {
        fprintf(stderr, "usage: %s <family> <matrix.mtx> <N> <S> <C> <dump_dir>\n", argv[0]);
        return 2;
    }
    string family = argv[1], mtx = argv[2];
    int n = atoi(argv[3]), s = atoi(argv[4]), c = atoi(argv[5]);
    g_dump_dir = argv[6];
    set_config("DENSE_MATRIX_SIZE", n); // token_test.cc:1630
    g_set = create_init_metadata_set_from_file(mtx, "");
    if (family.compare(0, 4, "ops=") == 0) return run_ops(family.substr(4));
    if (family == "load")
    {
        dump_all();
        return 0;
    }
    if (family == "thread_total") test_spmm_thread_total(g_set, mtx, s, c);
    else if (family == "warp_total") test_spmm_warp_total(g_set, mtx, s, c);
    else if (family == "warp_bitmap") test_spmm_warp_bitmap(g_set, mtx, s, c);
    else if (family == "thread_bit_map") test_spmm_thread_bit_map(g_set, mtx, s, c);
    else if (family == "warp_segment") test_spmm_warp_segment(g_set, mtx, s, c);
    else if (family == "block_total") test_spmm_block_total(g_set, mtx, s, c);
    else if (family == "block_bitmap") test_spmm_block_bitmap(g_set, mtx, s, c);
    else { fprintf(stderr, "unknown family %s\n", family.c_str()); return 2; }
    return 0;
}
