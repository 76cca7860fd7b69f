// main.cpp -- `curvix_main -c <settings.yaml>`: the reference's command line and pass loop (cmd/main.cpp:36-174) over the
// drop-in EdgeSketch_Core.  Extra, optional flags (none changes results):
//   --outputs DIR       where 3D_edges_*.txt / 3D_tangents_*.txt go   (default ../../outputs, definitions.h:4)
//   --dataset-path DIR  overrides the settings file's Dataset_Path
//   --device N          CUDA device ordinal
//   --devices 0,1,..    several GPUs: every pass sharded by view-1 edge range (balanced by hypothesis rows), results gathered over NCCL
//   --edge-cache DIR    binary cache of the parsed edge files
//   --max-passes N      overrides Max_Num_Of_3D_Edge_Sketch_Passes
//   --schedule FILE     one line per pass: what was run (views, hypotheses, validations, pairs, claimed ratio, timings)
//   --keep-outputs      do not `rm -f` the 3D_edges_* / 3D_tangents_* files of the outputs folder first
//   --no-consolidate    stop after the sketch passes (skip EdgeMapping::consolidate_3D_edges, cmd/main.cpp:157-163)
//   --support-links FILE     dump the 3-D edge -> supporting 2-D edges table (the input of the consolidation)
//   --consolidate-links FILE skip the sketch: read such a dump and run only the consolidation on it
// After the passes the 3-D edges are grouped into curves as in the reference (cmd/main.cpp:157-163): links / prunings / alignment on the
// device (cvx_graph_*), curve tracing on the host; final_curves.txt and the intermediate files go to the outputs folder.
#include <sys/stat.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <memory>
#include <string>

#include "EdgeSketch_Core.hpp"

static void usage() {
    std::cout << "Usage: ./curvix_main [flag] [argument]\n\noptions:\n  -h, --help         show this help message and exit\n"
                 "  -c, --config_file  path to the 3D Curvix setting file\n";   // definitions.h:69-72
}

int main(int argc, char** argv) {
    std::string config, schedule_path, links_path, consolidate_from;
    bool consolidate = true;
    curvix_host::HostOptions opt;
    std::string dataset_override;
    int max_passes = -1;
    bool keep_outputs = false;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto next = [&]() -> std::string { if (i + 1 >= argc) { usage(); std::exit(0); } return argv[++i]; };
        if (a == "-h" || a == "--help") { usage(); return 0; }
        else if (a == "-c" || a == "--config_file") config = next();
        else if (a == "--outputs") opt.outputs_dir = next();
        else if (a == "--dataset-path") dataset_override = next();
        else if (a == "--device") opt.device = std::atoi(next().c_str());
        else if (a == "--devices") {   // 0,1,2,...: every pass sharded by view-1 edge range over these GPUs
            const std::string l = next();
            size_t b = 0;
            while (b <= l.size()) { const size_t e = l.find(',', b); opt.devices.push_back(std::atoi(l.substr(b, e == std::string::npos ? e : e - b).c_str())); if (e == std::string::npos) break; b = e + 1; }
        }
        else if (a == "--edge-cache") opt.edge_cache_dir = next();
        else if (a == "--max-passes") max_passes = std::atoi(next().c_str());
        else if (a == "--schedule") schedule_path = next();
        else if (a == "--support-links") links_path = next();
        else if (a == "--consolidate-links") consolidate_from = next();
        else if (a == "--no-consolidate") consolidate = false;
        else if (a == "--keep-outputs") keep_outputs = true;
        else if (a == "--refined") opt.use_refined_poses = std::atoi(next().c_str());
        else if (a == "--quiet") opt.quiet = true;
        else { std::cerr << "ERROR: Invalid input arguments. See the following usage:" << std::endl; usage(); return 0; }
    }
    if (config.empty()) { std::cerr << "ERROR: No setting file is given!" << std::endl; usage(); return 0; }

    // outputs folder (cmd/main.cpp:78-98)
    struct stat st;
    if (stat(opt.outputs_dir.c_str(), &st) != 0) {
        const std::string cmd = "mkdir -p '" + opt.outputs_dir + "'";
        if (std::system(cmd.c_str()) != 0) { std::cerr << "ERROR: Failed to create the outputs folder " << opt.outputs_dir << std::endl; return 0; }
    }
    if (!keep_outputs) {
        const std::string cmd = "rm -f '" + opt.outputs_dir + "'/3D_edges_*.txt '" + opt.outputs_dir + "'/3D_tangents_*.txt";
        if (std::system(cmd.c_str()) != 0) std::cerr << "ERROR: cannot remove old files under " << opt.outputs_dir << std::endl;
    }

    curvix_host::FlatYaml settings;
    try {
        settings = curvix_host::FlatYaml::load(config);
        if (!dataset_override.empty()) settings.set("Dataset_Path", dataset_override);
    } catch (const std::exception& e) {
        std::cerr << "Exception: " << e.what() << std::endl;
        return 0;
    }

    std::unique_ptr<EdgeSketch_Core> core;
    try {
        core.reset(new EdgeSketch_Core(settings, opt));
    } catch (const std::exception& e) {
        std::cerr << "Exception: " << e.what() << std::endl;
        return 0;
    }
    EdgeSketch_Core& MWV_Edge_Rec = *core;
    if (max_passes >= 0) MWV_Edge_Rec.Max_3D_Edge_Sketch_Passes = max_passes;
    MWV_Edge_Rec.Read_Camera_Data();
    std::shared_ptr<EdgeMapping> edgeMapping = std::make_shared<EdgeMapping>();
    FILE* sched = schedule_path.empty() ? nullptr : std::fopen(schedule_path.c_str(), "w");

    int edge_sketch_pass_count = 0;
    if (!consolidate_from.empty()) {
        if (!edgeMapping->read_support_links(consolidate_from)) { std::cerr << "ERROR: cannot read " << consolidate_from << std::endl; return 1; }
        MWV_Edge_Rec.Max_3D_Edge_Sketch_Passes = 0;
    }
    while (edge_sketch_pass_count < MWV_Edge_Rec.Max_3D_Edge_Sketch_Passes) {   // cmd/main.cpp:123-155
        const int v1 = MWV_Edge_Rec.hyp01_view_indx, v2 = MWV_Edge_Rec.hyp02_view_indx;
        if (!opt.quiet) std::cout << "Selected views for hypotheses are " << v1 << " and " << v2 << std::endl;
        MWV_Edge_Rec.Set_Hypothesis_Views_Camera();
        MWV_Edge_Rec.Read_Edgels_Data();
        MWV_Edge_Rec.Set_Hypothesis_Views_Edgels();
        MWV_Edge_Rec.Run_3D_Edge_Sketch();
        MWV_Edge_Rec.Finalize_Edge_Pairs_and_Reconstruct_3D_Edges(edgeMapping);
        MWV_Edge_Rec.Stack_3D_Edges();
        MWV_Edge_Rec.Calculate_Edge_Support_Ratios_And_Select_Next_Views(edgeMapping);
        if (sched) {
            const cvx_counts& c = MWV_Edge_Rec.last_counts;
            std::fprintf(sched, "%d %d %d %lld %lld %lld %lld %.17g %.6f %.6f %.6f\n", edge_sketch_pass_count, v1, v2, (long long)c.n_hypotheses,
                         (long long)c.n_validations, (long long)c.n_pairs, (long long)MWV_Edge_Rec.Gamma1s.rows, MWV_Edge_Rec.avg_ratio,
                         MWV_Edge_Rec.pair_edges_time, MWV_Edge_Rec.finalize_edge_pair_time, MWV_Edge_Rec.find_next_hypothesis_view_time);
            std::fflush(sched);
        }
        MWV_Edge_Rec.Clear_Data();
        edge_sketch_pass_count++;
        if (MWV_Edge_Rec.enable_aborting_3D_edge_sketch) break;
    }
    if (sched) std::fclose(sched);
    if (!links_path.empty()) edgeMapping->write_support_links(links_path);

    if (consolidate) {   // cmd/main.cpp:157-163
        if (!opt.quiet) std::cout << "Start to merge edges from 3D-2D edge links and 2D-2D edge links ..." << std::endl;
        edgeMapping->Setup_Data_Parameters(settings);
        if (!edgeMapping->consolidate_3D_edges(MWV_Edge_Rec.gpu(), MWV_Edge_Rec.All_R, MWV_Edge_Rec.All_T, MWV_Edge_Rec.K, MWV_Edge_Rec.Num_Of_Total_Imgs,
                                               opt.outputs_dir, true, opt.quiet)) {
            std::cerr << "ERROR: " << edgeMapping->last_error << std::endl;
            return 1;
        }
        const EdgeMapping::ConsolidationReport& r = edgeMapping->report;
        std::cout << "     - (untimed in the reference) consolidation: " << r.total_s << " (s): " << r.nodes << " 3D edges, " << r.records << " 3D-2D links, "
                  << r.curvelets << " curvelets -> " << r.links << " / " << r.links_after_3d << " / " << r.links_after_projection
                  << " graph links (built / pruned in 3D / pruned by projection); graph " << r.graph_s << " s, " << edgeMapping->num_of_iterations
                  << " alignment iterations " << r.align_s << " s (kernels " << r.align_kernel_ms << " ms), curves " << r.curves_s << " s -> "
                  << r.curves_from_graph << " traced, " << r.final_curves << " final curves" << std::endl;
    }

    const double total_time = MWV_Edge_Rec.pair_edges_time + MWV_Edge_Rec.finalize_edge_pair_time + MWV_Edge_Rec.find_next_hypothesis_view_time;
    std::cout << "Total computation time: " << total_time << " (s)" << std::endl;   // cmd/main.cpp:165-170
    std::cout << "     - Time for Pairing up Edges:              " << MWV_Edge_Rec.pair_edges_time << " (s)" << std::endl;
    std::cout << "     - Time for Finalizing Edge Pairs:         " << MWV_Edge_Rec.finalize_edge_pair_time << " (s)" << std::endl;
    std::cout << "     - Time for Finding Next Hypothesis Views: " << MWV_Edge_Rec.find_next_hypothesis_view_time << " (s)" << std::endl;
    std::cout << "     - (untimed in the reference) Stacking 3D edges + writing the two files: " << MWV_Edge_Rec.stack_time << " (s)" << std::endl;
    std::cout << "     - (untimed in the reference) edge file reading + upload: " << MWV_Edge_Rec.read_edges_time << " (s)" << std::endl;
    std::cout << "     - passes: " << edge_sketch_pass_count << ", 3D edges: " << MWV_Edge_Rec.all_3D_Edges.rows << ", 3D-2D links: "
              << edgeMapping->num_support_links() << std::endl;
    std::cout << "3D Curvix is Finished!" << std::endl;
    return 0;
}
