/* TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * Driver that runs the REFERENCE's own model code (compiled from /root/reference/src/model/*.cpp
 * where it lies, see oracle/Makefile) on a flat network snapshot and dumps per-step state.
 * It replaces exactly the three reference files that cannot be built offline:
 *   - src/ModelRun.cpp          (includes NetCDF output)        -> ModelRun members below,
 *                                                                  same phase order as ModelRun.cpp:153-229
 *   - src/input/ModelInitializer.cpp (NetCDF + yaml-cpp input)  -> class ModelInitializer below, calling the
 *                                                                  same model mutators in the same order as
 *                                                                  ModelInitializer.cpp:73-213,689-700,1081-1085
 *   - src/scenario/Scenario.cpp (settings tree)                 -> forcing table applied with the rule of
 *                                                                  Scenario.cpp:161-174
 * No arithmetic of the hot path lives here: every number dumped is computed by reference code.
 */
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <limits>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "ModelRun.h"
#include "acclimate.h"
#include "accio.h"
#include "model/BusinessConnection.h"
#include "model/CapacityManager.h"
#include "model/Consumer.h"
#include "model/EconomicAgent.h"
#include "model/Firm.h"
#include "model/GeoConnection.h"
#include "model/GeoRoute.h"
#include "model/Model.h"
#include "model/PurchasingManager.h"
#include "model/Region.h"
#include "model/SalesManager.h"
#include "model/Sector.h"
#include "model/Storage.h"
#include "model/TransportChainLink.h"
#include "nlopt.h"
#include "openmp.h"

#include <unordered_set>

#ifndef ORACLE_NLOPT_KRAFT
// provider "shim" only (the product's solver behind NLopt's entry points): the one definition of the transport penalty's
// kink and the solver-group rule.  The provider "kraft" build (make kraft) includes nothing of the product.
#include "../../acclimate_b200/csrc/purchase_core.cuh"
#endif

// index of each entry in the "model_f64" array of the network file (acclimate_b200/netfile.py)
enum ModelParamIdx {
    MP_TRANSPORT_PENALTY_SMALL = 0,
    MP_TRANSPORT_PENALTY_LARGE,
    MP_MIN_STORAGE,
    MP_CHEAPEST_PRICE_RANGE_WIDTH,  // < 0: "auto"
    MP_OPTIMIZATION_MAXITER,
    MP_OPTIMIZATION_PRECISION_ADJUSTMENT,
    MP_QUADRATIC_TRANSPORT_PENALTY,
    MP_RELATIVE_TRANSPORT_PENALTY,
    MP_MAX_DECREASE_RESERVATION_PRICE_LIMITED_BY_MARKUP,
    MP_DEVIATION_PENALTY,
    MP_CHEAPEST_PRICE_RANGE_PRESERVE_SELLER_PRICE,
    MP_NO_SELF_SUPPLY,
    MP_UTILITARIAN,
    MP_DELTA_T,
    MP_NAIVE_EXPECTATIONS,
    MP_ALWAYS_EXTEND_EXPECTED_DEMAND_CURVE,
    MP_COUNT
};
// per-sector parameters in "sector_f64" (row-major, SP_COUNT per sector)
enum SectorParamIdx {
    SP_UPPER_STORAGE_LIMIT = 0,
    SP_INITIAL_STORAGE_FILL_FACTOR,
    SP_TRANSPORT_TYPE,  // 0 aviation, 1 immediate, 2 roadsea
    SP_SUPPLY_ELASTICITY,
    SP_PRICE_INCREASE_PRODUCTION_EXTENSION,
    SP_ESTIMATED_PRICE_INCREASE_PRODUCTION_EXTENSION,
    SP_INITIAL_MARKUP,
    SP_TARGET_STORAGE_REFILL_TIME,
    SP_TARGET_STORAGE_WITHDRAW_TIME,
    SP_COUNT
};

struct HarnessConfig {
    std::map<std::string, accio::Array> net;
    int steps = 1;
    std::string dump_path;
    int dump_every = 1;
    bool dump_edges = true;
    std::string ingest_out;         // --ingest <out.acc>: candidate flow table -> threshold filter -> clean_network; dumps the survivors, no run
    bool dump_observables = false;  // --observe: every observable of the reference's observe<>() visitors, shaped like ArrayOutput's data
    const std::vector<std::int64_t>& I(const char* n) const { return net.at(n).i; }
    const std::vector<double>& F(const char* n) const { return net.at(n).f; }
};

namespace settings {
class SettingsNode {  // stand-in: the harness passes its own config through the reference's constructor signature
  public:
    const HarnessConfig* cfg;
};
}  // namespace settings

struct EventRec {
    std::int64_t step, type, a, b;
    double value;
};
static std::vector<EventRec> g_events;
static std::unordered_set<const void*> g_purchasing_managers;  // which nlopt user data are PurchasingManagers
static std::int64_t g_step = 0;

namespace acclimate {

class Scenario {};  // complete types for ModelRun's unique_ptr members (unused)
class Output {};

const std::array<options::Option, 10> options::options = {{{"BASED_ON_INT", true},
                                                           {"BANKERS_ROUNDING", true},
                                                           {"CHECKPOINTING", false},
                                                           {"CLEANUP_INFO", false},
                                                           {"FATAL_FLOATING_POINT_EXCEPTIONS", false},
                                                           {"FLOATING_POINT_EXCEPTIONS", false},
                                                           {"OPTIMIZATION_PROBLEMS_FATAL", false},
                                                           {"OPTIMIZATION_WARNINGS", false},
                                                           {"STRICT_MIN_DERIVATIVE", false},
                                                           {"USE_MIN_PASSAGE_IN_EXPECTATION", true}}};

// ---- network construction (follows src/input/ModelInitializer.cpp) ------------------------------
class ModelInitializer {
  public:
    Model* model;
    const HarnessConfig& cfg;
    std::vector<EconomicAgent*> agents;
    std::vector<std::shared_ptr<BusinessConnection>> edges;  // file order
    std::vector<Storage*> storages;                          // creation order
    std::vector<std::shared_ptr<GeoConnection>> connections; // unordered region pairs, (i, j<i) order
    std::vector<std::int64_t> conn_a, conn_b;

    ModelInitializer(Model* m, const HarnessConfig& c) : model(m), cfg(c) {}

    void run() {
        const auto& meta = cfg.I("meta");
        const std::size_t n_sectors = meta[0], n_regions = meta[1], n_agents = meta[2], n_edges = meta[3];
        const auto& mp = cfg.F("model_f64");
        const auto& sp = cfg.F("sector_f64");
        // ModelInitializer.cpp:58-62
        model->set_delta_t(Time(mp[MP_DELTA_T]));
        model->no_self_supply(mp[MP_NO_SELF_SUPPLY] != 0.0);
        // pre_initialize, ModelInitializer.cpp:1015-1079 (defaults of the reference for everything not in the file)
        auto& P = model->parameters_writable();
        P.transport_penalty_small = Price(mp[MP_TRANSPORT_PENALTY_SMALL]);
        P.transport_penalty_large = Price(mp[MP_TRANSPORT_PENALTY_LARGE]);
        P.optimization_maxiter = int(mp[MP_OPTIMIZATION_MAXITER]);
        P.optimization_timeout = 3600;
        P.quadratic_transport_penalty = mp[MP_QUADRATIC_TRANSPORT_PENALTY] != 0.0;
        P.maximal_decrease_reservation_price_limited_by_markup = mp[MP_MAX_DECREASE_RESERVATION_PRICE_LIMITED_BY_MARKUP] != 0.0;
        P.always_extend_expected_demand_curve = mp.size() > MP_ALWAYS_EXTEND_EXPECTED_DEMAND_CURVE && mp[MP_ALWAYS_EXTEND_EXPECTED_DEMAND_CURVE] != 0.0;
        P.naive_expectations = !(mp.size() > MP_NAIVE_EXPECTATIONS) || mp[MP_NAIVE_EXPECTATIONS] != 0.0;
        P.deviation_penalty = mp[MP_DEVIATION_PENALTY] != 0.0;
        P.min_storage = mp[MP_MIN_STORAGE];
        P.cheapest_price_range_preserve_seller_price = mp[MP_CHEAPEST_PRICE_RANGE_PRESERVE_SELLER_PRICE] != 0.0;
        P.cheapest_price_range_generic_size = mp[MP_CHEAPEST_PRICE_RANGE_WIDTH] < 0.0;
        if (!P.cheapest_price_range_generic_size) P.cheapest_price_range_width = Price(mp[MP_CHEAPEST_PRICE_RANGE_WIDTH]);
        P.relative_transport_penalty = mp[MP_RELATIVE_TRANSPORT_PENALTY] != 0.0;
        P.respect_markup_in_production_extension = false;  // never assigned by the reference (parameters.h:46); pinned
        P.optimization_algorithm = NLOPT_LD_SLSQP;
        P.utility_optimization_algorithm = NLOPT_LD_SLSQP;
        P.global_optimization_algorithm = NLOPT_GN_CRS2_LM;
        P.lagrangian_algorithm = NLOPT_AUGLAG;
        P.global_utility_optimization_algorithm = NLOPT_G_MLSL_LDS;
        P.global_purchasing_optimization = false;
        P.local_purchasing_optimization = true;
        P.optimization_restart_baseline = false;
        P.start_purchasing_at_baseline = false;
        P.purchasing_halfway_baseline = false;
        P.global_utility_optimization = false;
        P.budget_inequality_constrained = false;
        P.elastic_budget = false;
        P.global_utility_optimization_random_points = 64;
        P.utility_optimization_maxiter = P.optimization_maxiter;
        P.utility_optimization_timeout = P.optimization_timeout;
        P.global_optimization_maxiter = P.optimization_maxiter;
        P.global_optimization_timeout = P.optimization_timeout;
        P.global_utility_optimization_maxiter = P.optimization_maxiter;
        P.global_utility_optimization_timeout = P.optimization_timeout;
        P.optimization_precision_adjustment = int(mp[MP_OPTIMIZATION_PRECISION_ADJUSTMENT]);
        P.utility_optimization_precision_adjustment = 1;
        P.global_optimization_precision_adjustment = 1;
        P.global_utility_optimization_precision_adjustment = 1;

        // sectors: add_sector, ModelInitializer.cpp:159-185
        for (std::size_t s = 0; s < n_sectors; ++s) {
            const double* q = &sp[s * SP_COUNT];
            char name[32];
            if (s + 1 == n_sectors) std::snprintf(name, sizeof(name), "FCON");
            else std::snprintf(name, sizeof(name), "SEC%02zu", s);
            const int tt = int(q[SP_TRANSPORT_TYPE]);
            auto* sector = model->sectors.add(model, id_t(name), Ratio(q[SP_UPPER_STORAGE_LIMIT]), q[SP_INITIAL_STORAGE_FILL_FACTOR] * model->delta_t(),
                                              tt == 0   ? Sector::transport_type_t::AVIATION
                                              : tt == 1 ? Sector::transport_type_t::IMMEDIATE
                                                        : Sector::transport_type_t::ROADSEA);
            sector->parameters_writable().supply_elasticity = q[SP_SUPPLY_ELASTICITY];
            sector->parameters_writable().price_increase_production_extension = Price(q[SP_PRICE_INCREASE_PRODUCTION_EXTENSION]);
            sector->parameters_writable().estimated_price_increase_production_extension = Price(q[SP_ESTIMATED_PRICE_INCREASE_PRODUCTION_EXTENSION]);
            sector->parameters_writable().initial_markup = Price(q[SP_INITIAL_MARKUP]);
            sector->parameters_writable().target_storage_refill_time = q[SP_TARGET_STORAGE_REFILL_TIME] * model->delta_t();
            sector->parameters_writable().target_storage_withdraw_time = q[SP_TARGET_STORAGE_WITHDRAW_TIME] * model->delta_t();
        }
        for (std::size_t r = 0; r < n_regions; ++r) {
            char name[32];
            std::snprintf(name, sizeof(name), "RG%03zu", r);
            model->regions.add(model, id_t(name));
        }
        // agents: add_firm / add_consumer, ModelInitializer.cpp:94-142
        const auto& at = cfg.I("agent_type");
        const auto& as = cfg.I("agent_sector");
        const auto& ar = cfg.I("agent_region");
        const auto& ab = cfg.F("agent_beta");
        const bool utilitarian = mp[MP_UTILITARIAN] != 0.0;
        for (std::size_t a = 0; a < n_agents; ++a) {
            Sector* sector = model->sectors[as[a]];
            Region* region = model->regions[ar[a]];
            const std::string name = sector->name() + ":" + region->name();
            if (at[a] == 1) {
                auto* firm = model->economic_agents.add<Firm>(id_t(name), sector, region, Ratio(ab[a]));
                sector->firms.add(firm);
                region->economic_agents.add(firm);
                agents.push_back(firm);
            } else {
                // consumer_baskets / inter_basket_substitution_coefficient, ModelInitializer.cpp:118-139.  Snapshot keys
                // basket_of_sector / basket_coeff / inter_basket_coeff; without them: one basket of all supplying sectors
                // (substitution 2.0 inside, 1.5 between baskets).  Sectors of a basket are listed in ascending order.
                std::vector<std::pair<std::vector<Sector*>, FloatType>> baskets;
                FloatType inter = 1.5;
                if (cfg.net.count("basket_of_sector") != 0) {
                    const auto& bos = cfg.I("basket_of_sector");
                    const auto& bco = cfg.F("basket_coeff");
                    inter = cfg.F("inter_basket_coeff").at(0);
                    for (std::size_t b = 0; b < bco.size(); ++b) {
                        std::vector<Sector*> members;
                        for (std::size_t s = 0; s < n_sectors; ++s) if (bos.at(s) == std::int64_t(b)) members.push_back(model->sectors[s]);
                        baskets.emplace_back(members, bco[b]);
                    }
                } else {
                    std::vector<Sector*> all;
                    for (std::size_t s = 0; s + 1 < n_sectors; ++s) all.push_back(model->sectors[s]);
                    baskets.emplace_back(all, 2.0);
                }
                auto* consumer = model->economic_agents.add<Consumer>(id_t(name), region, inter, baskets, utilitarian);
                region->economic_agents.add(consumer);
                agents.push_back(consumer);
            }
        }
        // transport: one simple connection per unordered region pair (create_simple_transport_connection,
        // ModelInitializer.cpp:689-700; loop order of build_transport_network "const", :985-992)
        if (cfg.net.count("entity_delay") != 0) {
            // multi-hop routes ("network" transport, ModelInitializer.cpp:575-607): snapshot keys entity_delay[n_entities],
            // route_off[2 R R + 1] / route_entity[]: the geo entities a shipment passes between two regions, for
            // aviation (k = 0) and road/sea (k = 1) sectors, at index (k R + from) R + to.  Every entity is created as a
            // GeoConnection: a transport chain link only reads its delay and forcing (TransportChainLink.cpp:40-63).
            const auto& ed = cfg.I("entity_delay");
            const auto& ro = cfg.I("route_off");
            const auto& re = cfg.I("route_entity");
            for (std::size_t k = 0; k < ed.size(); ++k) {
                connections.push_back(std::make_shared<GeoConnection>(model, TransportDelay(ed[k]), GeoConnection::type_t::UNSPECIFIED, model->regions[0],
                                                                      model->regions[n_regions > 1 ? 1 : 0]));
            }
            const Sector::transport_type_t kinds[2] = {Sector::transport_type_t::AVIATION, Sector::transport_type_t::ROADSEA};
            for (std::size_t k = 0; k < 2; ++k)
                for (std::size_t i = 0; i < n_regions; ++i)
                    for (std::size_t j = 0; j < n_regions; ++j) {
                        if (i == j) continue;
                        const std::size_t r = (k * n_regions + i) * n_regions + j;
                        GeoRoute route;
                        for (std::int64_t q = ro.at(r); q < ro.at(r + 1); ++q) route.path.add(connections.at(std::size_t(re.at(std::size_t(q)))).get());
                        if (route.path.size() == 0) throw std::runtime_error("empty route in the snapshot");
                        model->regions[i]->routes.emplace(std::make_pair(IndexType(model->regions[j]->id.index()), kinds[k]), route);
                    }
        } else {
        const auto& delay = cfg.I("delay");
        for (std::size_t i = 0; i < n_regions; ++i) {
            for (std::size_t j = 0; j < i; ++j) {
                Region* from = model->regions[i];
                Region* to = model->regions[j];
                auto inf = std::make_shared<GeoConnection>(model, TransportDelay(delay[i * n_regions + j]), GeoConnection::type_t::UNSPECIFIED, from, to);
                from->connections.push_back(inf);
                to->connections.push_back(inf);
                GeoRoute route;
                route.path.add(inf.get());
                from->routes.emplace(std::make_pair(IndexType(to->id.index()), Sector::transport_type_t::AVIATION), route);
                from->routes.emplace(std::make_pair(IndexType(to->id.index()), Sector::transport_type_t::ROADSEA), route);
                to->routes.emplace(std::make_pair(IndexType(from->id.index()), Sector::transport_type_t::AVIATION), route);
                to->routes.emplace(std::make_pair(IndexType(from->id.index()), Sector::transport_type_t::ROADSEA), route);
                connections.push_back(inf);
                conn_a.push_back(i);
                conn_b.push_back(j);
            }
        }
        }
        // connections: initialize_connection, ModelInitializer.cpp:187-213
        std::vector<std::int64_t> t_from, t_to, t_zq;
        const bool ingest = !cfg.ingest_out.empty();
        if (ingest) {
            // the dense branch of build_agent_network_from_table, ModelInitializer.cpp:832-834,947-965: annual flows as float,
            // kept when  *d > flow_threshold && flow > daily_flow_threshold  with flow = round(FlowQuantity(*d) * time_factor)
            const auto& table = cfg.F("flows_f32");  // values exactly representable as float, row = source agent
            const FloatType flow_threshold = cfg.F("threshold").at(0);
            const FloatType time_factor = cfg.F("time_factor").at(0);
            const FlowQuantity daily_flow_threshold = round(FlowQuantity(flow_threshold * time_factor));
            if (table.size() != n_agents * n_agents) throw std::runtime_error("flows_f32 must hold n_agents x n_agents values");
            for (std::size_t i = 0; i < n_agents; ++i) {
                if (!agents[i]->is_firm()) continue;
                for (std::size_t j = 0; j < n_agents; ++j) {
                    const float d = float(table[i * n_agents + j]);
                    const FlowQuantity flow = round(FlowQuantity(d) * time_factor);
                    if (d > flow_threshold && flow > daily_flow_threshold) {
                        if (model->no_self_supply() && i == j) continue;  // initialize_connection, :188-190
                        t_from.push_back(std::int64_t(i)); t_to.push_back(std::int64_t(j));
                        t_zq.push_back(iround(to_float(flow) / FlowQuantity::precision));
                    }
                }
            }
        }
        const auto& ef = ingest ? t_from : cfg.I("edge_from");
        const auto& et = ingest ? t_to : cfg.I("edge_to");
        const auto& ez = ingest ? t_zq : cfg.I("edge_zq");
        const auto& cel = cfg.F("consumer_elasticity");
        const std::size_t n_edges_used = ingest ? t_from.size() : n_edges;
        edges.reserve(n_edges_used);
        for (std::size_t e = 0; e < n_edges_used; ++e) {
            Firm* firm_from = agents[ef[e]]->as_firm();
            EconomicAgent* to = agents[et[e]];
            if (model->no_self_supply() && static_cast<void*>(firm_from) == static_cast<void*>(to)) {
                throw std::runtime_error("network file contains a self-supply edge although no_self_supply is set");
            }
            Sector* sector_from = firm_from->sector;
            auto* input_storage = to->input_storages.find(sector_from->name() + "->" + to->name());
            if (input_storage == nullptr) {
                input_storage = to->input_storages.add(sector_from, to);
                if (to->is_consumer()) input_storage->parameters_writable().consumption_price_elasticity = cel[sector_from->id.index()];
                storages.push_back(input_storage);
            }
            const Flow flow = Flow(FlowQuantity(double(ez[e]) * FlowQuantity::precision), Price(1.0));
            if (iround(to_float(flow.get_quantity()) / FlowQuantity::precision) != ez[e]) throw std::runtime_error("flow quantisation mismatch");
            input_storage->add_initial_flow_Z_star(flow);
            firm_from->add_initial_production_X_star(flow);
            auto bc = std::make_shared<BusinessConnection>(input_storage->purchasing_manager.get(), firm_from->sales_manager.get(), flow);
            firm_from->sales_manager->business_connections.emplace_back(bc);
            input_storage->purchasing_manager->business_connections.emplace_back(bc);
            if (static_cast<void*>(firm_from) == static_cast<void*>(to)) firm_from->self_supply_connection(bc);
            edges.push_back(bc);
        }
        std::size_t n_max = 0;
        for (auto* st : storages) {
            g_purchasing_managers.insert(st->purchasing_manager.get());
            n_max = std::max(n_max, st->purchasing_manager->business_connections.size());
        }
#ifndef ORACLE_NLOPT_KRAFT
        nlopt_provider_set_group(accsqp_group_for(int(n_max)));  // the engine's rule for its solver group (csrc/sqp_core.h)
#endif
        if (ingest) {
            clean_network();
            dump_survivors(ef, et, ez);
            return;
        }
        check_clean();
        // post_initialize, ModelInitializer.cpp:1081-1085
        for (auto& agent : model->economic_agents) agent->initialize();
    }

    // clean_network, ModelInitializer.cpp:215-299, restated around the reference's own mutators (Storage::subtract_initial_flow_Z_star,
    // PurchasingManager / SalesManager::remove_business_connection, Firm::subtract_initial_production_X_star): sweeps over the
    // agents in order; a firm without value added, buyers or inputs (a lone self-supply connection does not count) is taken out
    // at once -- its buyers' storages and its suppliers' production shrink before the next agent is looked at -- and the
    // sweeps repeat until one removes nothing.
    std::size_t clean_sweeps = 0;
    void clean_network() {
        for (;;) {
            ++clean_sweeps;
            std::vector<EconomicAgent*> to_remove;
            for (auto& ea : model->economic_agents) {
                if (ea->is_firm()) {
                    Firm* firm = ea->as_firm();
                    FlowQuantity input(0.0);
                    for (const auto& is : firm->input_storages) input = input + is->initial_input_flow_I_star().get_quantity();
                    const FloatType value_added = to_float(firm->initial_production_X_star().get_quantity() - input);
                    const bool self = firm->self_supply_connection() != nullptr;
                    const auto& out = firm->sales_manager->business_connections;
                    if (value_added <= 0.0 || out.empty() || (out.size() == 1 && self) || firm->input_storages.empty() || (firm->input_storages.size() == 1 && self)) {
                        for (auto& bc : firm->sales_manager->business_connections) {            // :253-260: the buyers lose this input
                            if (bc->buyer == nullptr) throw std::runtime_error("Buyer invalid");
                            if (!bc->buyer->storage->subtract_initial_flow_Z_star(bc->initial_flow_Z_star())) bc->buyer->remove_business_connection(bc.get());
                        }
                        for (auto& storage : firm->input_storages) {                            // :263-271: the suppliers lose this buyer
                            for (auto& bc : storage->purchasing_manager->business_connections) {
                                if (bc->seller == nullptr) throw std::runtime_error("Seller invalid");
                                bc->seller->firm->subtract_initial_production_X_star(bc->initial_flow_Z_star());
                                bc->seller->remove_business_connection(bc.get());
                            }
                        }
                        firm->sector->firms.remove(firm);
                        firm->region->economic_agents.remove(ea.get());
                        to_remove.push_back(ea.get());
                    }
                } else if (ea->input_storages.empty()) {                                        // :279-286
                    ea->region->economic_agents.remove(ea.get());
                    to_remove.push_back(ea.get());
                }
            }
            if (to_remove.empty()) break;
            for (auto* gone : to_remove)
                for (auto& slot : agents) if (slot == gone) slot = nullptr;
            model->economic_agents.remove(to_remove);
        }
    }
    // what is left: agents by their index in the candidate list, connections that still link a living seller to a living storage
    void dump_survivors(const std::vector<std::int64_t>& ef, const std::vector<std::int64_t>& et, const std::vector<std::int64_t>& ez) {
        accio::Writer out(cfg.ingest_out);
        std::vector<std::int64_t> alive(agents.size(), 0), o_from, o_to, o_zq, xstar(agents.size(), 0);
        for (std::size_t a = 0; a < agents.size(); ++a) {
            if (agents[a] == nullptr) continue;
            alive[a] = 1;
            if (agents[a]->is_firm()) {
                xstar[a] = iround(to_float(agents[a]->as_firm()->initial_production_X_star().get_quantity()) / FlowQuantity::precision);
                for (const auto& bc : agents[a]->as_firm()->sales_manager->business_connections) {
                    const EconomicAgent* buyer = bc->buyer->storage->economic_agent;
                    std::size_t b = 0;
                    while (b < agents.size() && agents[b] != buyer) ++b;
                    o_from.push_back(std::int64_t(a)); o_to.push_back(std::int64_t(b));
                    o_zq.push_back(iround(to_float(bc->initial_flow_Z_star().get_quantity()) / FlowQuantity::precision));
                }
            }
        }
        (void)ef; (void)et; (void)ez;
        out.put("agent_alive", alive);
        out.put("edge_from", o_from); out.put("edge_to", o_to); out.put("edge_zq", o_zq);
        out.put("firm_Xstar_q", xstar);
        out.put("sweeps", std::vector<std::int64_t>{std::int64_t(clean_sweeps)});
    }

    // clean_network (ModelInitializer.cpp:215-299) must be a no-op for a valid snapshot: the generator guarantees it.
    void check_clean() {
        for (auto& ea : model->economic_agents) {
            if (ea->is_firm()) {
                Firm* firm = ea->as_firm();
                FlowQuantity input(0.0);
                for (const auto& is : firm->input_storages) input = input + is->initial_input_flow_I_star().get_quantity();
                const FloatType value_added = to_float(firm->initial_production_X_star().get_quantity() - input);
                if (value_added <= 0.0 || firm->sales_manager->business_connections.empty() || firm->input_storages.empty()) {
                    throw std::runtime_error("network snapshot is not clean: " + firm->name() + " would be removed by clean_network");
                }
            } else if (ea->input_storages.empty()) {
                throw std::runtime_error("network snapshot is not clean: consumer without inputs");
            }
        }
    }
};

}  // namespace acclimate

// ---- kinks of the purchasing objective for the NLopt shim (nlopt_shim.cpp: nlopt_shim_kinks) ----
// The reference's callbacks hand NLopt a gradient that jumps at D_r == target (PurchasingManager.cpp:388-396,432-450).
// The engine's solver is told where (csrc/purchase_core.cuh PMWork::W); this hook tells the shim the same, from the
// reference's own objects and with the same formula (acc::pm_kink_slopes), so both run identical arithmetic.
#ifndef ORACLE_NLOPT_KRAFT
extern "C" int nlopt_shim_kinks(void* f_data, unsigned n, double* scale, double* target, double* sl, double* sr, double* mid) {
    using namespace acclimate;
    if (g_purchasing_managers.count(f_data) == 0) return 0;
    const auto* pm = static_cast<const PurchasingManager*>(f_data);
    if (pm->purchasing_connections.size() != n) return 0;
    const auto& P = pm->model()->parameters();
    acc::PMParams pp;
    pp.Uf = to_float(pm->storage->initial_used_flow_U_star().get_quantity());
    pp.markup = to_float(pm->storage->sector->parameters().initial_markup);
    pp.pen_large = to_float(P.transport_penalty_large);
    pp.pen_small = to_float(P.transport_penalty_small);
    pp.quadratic = P.quadratic_transport_penalty; pp.relative = P.relative_transport_penalty;
    pp.limited_by_markup = P.maximal_decrease_reservation_price_limited_by_markup;
    for (unsigned r = 0; r < n; ++r) {
        const BusinessConnection* bc = pm->purchasing_connections[r];
        const double zf = to_float(bc->initial_flow_Z_star().get_quantity());
        const double t = P.deviation_penalty ? to_float(bc->last_demand_request_D(pm).get_quantity()) : zf;
        const acc::KinkSlopes ks = acc::pm_kink_slopes(pp, t);
        const double z_over_u = zf / pp.Uf;
        scale[r] = zf; target[r] = t;
        sl[r] = acc::pm_scaled_slope(ks.sl, z_over_u); sr[r] = acc::pm_scaled_slope(ks.sr, z_over_u); mid[r] = acc::pm_scaled_slope(ks.mid, z_over_u);
    }
    return 1;
}
#endif

// ---- ModelRun members (replacement for src/ModelRun.cpp) -------------------------------------------
namespace acclimate {

static ModelInitializer* g_init = nullptr;
static const HarnessConfig* g_cfg = nullptr;
static double g_phase_seconds = 0.0;

static inline std::int64_t qi(const FlowQuantity& q) { return iround(to_float(q) / FlowQuantity::precision); }
static inline std::int64_t qi(const Quantity& q) { return iround(to_float(q) / Quantity::precision); }
static inline std::int64_t pi(const Price& p) { return iround(to_float(p) / Price::precision); }

ModelRun::ModelRun(const settings::SettingsNode& settings) {
    step(IterationStep::INITIALIZATION);
    fesetround(FE_TONEAREST);  // ModelRun.cpp:80-82
    g_cfg = settings.cfg;
    start_time_m = Time(0.0);
    stop_time_m = Time(double(g_cfg->steps - 1) * g_cfg->F("model_f64")[MP_DELTA_T]);
    auto model = new Model(this);
    g_init = new ModelInitializer(model, *g_cfg);
    g_init->run();
    model_m.reset(model);
}

ModelRun::~ModelRun() {
    outputs_m.clear();
    model_m.reset();
}

bool ModelRun::done() const { return model_m->time() > stop_time_m; }

unsigned int ModelRun::thread_count() const { return openmp::get_thread_count(); }

std::string ModelRun::timeinfo() const { return std::to_string(time_m); }

static void record_event(EventType type, std::int64_t a, std::int64_t b, FloatType value) {
#pragma omp critical(harness_events)
    g_events.push_back(EventRec{g_step, std::int64_t(type), a, b, value});
}

void ModelRun::event(EventType type, const Sector* sector, const EconomicAgent* economic_agent, FloatType value) {
    record_event(type, std::int64_t(sector->id.index()), std::int64_t(economic_agent->id.index()), value);
}
void ModelRun::event(EventType type, const EconomicAgent* economic_agent, FloatType value) {
    record_event(type, std::int64_t(economic_agent->id.index()), -1, value);
}
void ModelRun::event(EventType type, const EconomicAgent* economic_agent_from, const EconomicAgent* economic_agent_to, FloatType value) {
    record_event(type, std::int64_t(economic_agent_from->id.index()), economic_agent_to ? std::int64_t(economic_agent_to->id.index()) : -1, value);
}

// forcing table applied with the rule of Scenario::iterate (src/scenario/Scenario.cpp:161-174)
static void apply_scenario(Model* model) {
    const auto it_f = g_cfg->net.find("forcing");
    if (it_f != g_cfg->net.end()) {
        const auto& f = it_f->second.f;
        for (std::size_t k = 0; k + 3 < f.size(); k += 4) {
            const Time from(f[k]), to(f[k + 1]);
            EconomicAgent* agent = g_init->agents[std::size_t(f[k + 2])];
            if (model->time() >= from && model->time() <= to) agent->set_forcing(f[k + 3]);
            else if (model->time() == to + model->delta_t()) agent->set_forcing(1.0);
        }
    }
    const auto it_p = g_cfg->net.find("passage");
    if (it_p != g_cfg->net.end()) {
        const auto& f = it_p->second.f;
        for (std::size_t k = 0; k + 3 < f.size(); k += 4) {
            const Time from(f[k]), to(f[k + 1]);
            GeoConnection* conn = g_init->connections[std::size_t(f[k + 2])].get();
            if (model->time() >= from && model->time() <= to) conn->set_forcing_nu(f[k + 3]);
            else if (model->time() == to + model->delta_t()) conn->set_forcing_nu(-1.);
        }
    }
}

// ---- the reference's OUTPUT SURFACE: every observable its observe<Observer, H>() visitors offer (include/model/*.h),
// collected the way ArrayOutput does (src/output/ArrayOutput.cpp:115-158: Flow / Stock split into <name>_quantity and
// <name>_value, everything through to_float) into arrays shaped like ArrayOutput::Variable::data of ONE timestep
// (ArrayOutput::iterate, :479-646): firms / consumers [agents], sectors, regions, storages [sectors][agents],
// flows [agent_from][agent_to]; cells nothing writes stay NaN (:170).
struct ObsDump {
    struct H {
        static constexpr const char* hash(const char* s) { return s; }
    };
    std::map<std::string, std::vector<double>> data;
    std::string group;
    std::size_t size = 0, index = 0;
    std::vector<double>& col(const std::string& name) {
        auto it = data.find(name);
        if (it == data.end()) it = data.emplace(name, std::vector<double>(size, std::numeric_limits<double>::quiet_NaN())).first;
        return it->second;
    }
    template<typename Function>
    auto set(const char* name, Function&& f) ->
        typename std::enable_if<!std::is_same<decltype(f()), Flow>::value && !std::is_same<decltype(f()), Stock>::value, bool>::type {
        col(group + ":" + name)[index] = to_float(f());
        return true;
    }
    template<typename Function>
    auto set(const char* name, Function&& f) ->
        typename std::enable_if<std::is_same<decltype(f()), Flow>::value || std::is_same<decltype(f()), Stock>::value, bool>::type {
        const auto& v = f();
        col(group + ":" + name + "_quantity")[index] = to_float(v.get_quantity());
        col(group + ":" + name + "_value")[index] = to_float(v.get_value());
        return true;
    }
    template<typename T>
    void collect(const T* v, std::size_t index_p) {
        index = index_p;
        v->template observe<ObsDump, H>(*this);
    }
    void begin(const char* g, std::size_t n) { group = g; size = n; }
};

static void dump_observables(accio::Writer& out, Model* model, const std::string& P) {
    ObsDump o;
    const auto& A = model->economic_agents;
    const std::size_t nA = A.size(), nS = model->sectors.size(), nR = model->regions.size();
    o.begin("firms", nA);
    for (std::size_t i = 0; i < nA; ++i) if (A[i]->is_firm()) o.collect(A[i]->as_firm(), i);
    o.begin("consumers", nA);
    for (std::size_t i = 0; i < nA; ++i) if (A[i]->is_consumer()) o.collect(A[i]->as_consumer(), i);
    o.begin("sectors", nS);
    for (std::size_t i = 0; i < nS; ++i) o.collect(model->sectors[i], i);
    o.begin("regions", nR);
    for (std::size_t i = 0; i < nR; ++i) o.collect(model->regions[i], i);
    o.begin("storages", nS * nA);
    for (std::size_t i = 0; i < nA; ++i)
        for (const auto& storage : A[i]->input_storages) o.collect(storage.get(), storage->sector->id.index() * nA + i);
    o.begin("flows", nA * nA);
    for (std::size_t i = 0; i < nA; ++i)
        if (A[i]->is_firm())
            for (const auto& bc : A[i]->as_firm()->sales_manager->business_connections) o.collect(bc.get(), i * nA + bc->buyer->storage->economic_agent->id.index());
    for (const auto& kv : o.data) out.put(P + "obs:" + kv.first, kv.second);
}

static void dump_step(accio::Writer& out, Model* model, std::int64_t t, bool with_edges) {
    char pre[32];
    std::snprintf(pre, sizeof(pre), "t%05ld/", long(t));
    const std::string P(pre);
    std::vector<std::int64_t> vi;
    std::vector<double> vf;
    auto put_i = [&](const char* name, std::size_t n, auto fn) {
        vi.resize(n);
        for (std::size_t k = 0; k < n; ++k) vi[k] = fn(k);
        out.put(P + name, vi);
    };
    auto put_f = [&](const char* name, std::size_t n, auto fn) {
        vf.resize(n);
        for (std::size_t k = 0; k < n; ++k) vf[k] = fn(k);
        out.put(P + name, vf);
    };
    const auto& E = g_init->edges;
    if (with_edges) {
        put_i("edge_D_q", E.size(), [&](std::size_t k) { return qi(E[k]->last_demand_request_D().get_quantity()); });
        put_i("edge_D_p", E.size(), [&](std::size_t k) { return pi(E[k]->last_demand_request_D().get_price()); });
        put_i("edge_ship_q", E.size(), [&](std::size_t k) { return qi(E[k]->last_shipment_Z().get_quantity()); });
        put_f("edge_ship_v", E.size(), [&](std::size_t k) { return to_float(E[k]->last_shipment_Z().get_value()); });
        put_i("edge_deliv_q", E.size(), [&](std::size_t k) { return qi(E[k]->last_delivery_Z().get_quantity()); });
        put_f("edge_deliv_v", E.size(), [&](std::size_t k) { return to_float(E[k]->last_delivery_Z().get_value()); });
        put_i("edge_deficit_q", E.size(), [&](std::size_t k) { return qi(E[k]->get_flow_deficit()); });
        put_i("edge_transit_q", E.size(), [&](std::size_t k) { return qi(E[k]->get_transport_flow().get_quantity()); });
        put_f("edge_transit_v", E.size(), [&](std::size_t k) { return to_float(E[k]->get_transport_flow().get_value()); });
    }
    const auto& S = g_init->storages;
    put_i("sto_S_q", S.size(), [&](std::size_t k) { return qi(S[k]->content_S().get_quantity()); });
    put_f("sto_S_v", S.size(), [&](std::size_t k) { return to_float(S[k]->content_S().get_value()); });
    put_i("sto_I_q", S.size(), [&](std::size_t k) { return qi(S[k]->last_input_flow_I().get_quantity()); });
    put_f("sto_I_v", S.size(), [&](std::size_t k) { return to_float(S[k]->last_input_flow_I().get_value()); });
    put_i("sto_Inext_q", S.size(), [&](std::size_t k) { return qi(S[k]->current_input_flow_I().get_quantity()); });  // register "other" == filled this step? see DESIGN
    put_i("sto_U_q", S.size(), [&](std::size_t k) { return qi(S[k]->used_flow_U().get_quantity()); });
    put_f("sto_U_v", S.size(), [&](std::size_t k) { return to_float(S[k]->used_flow_U().get_value()); });
    put_i("sto_Ut_q", S.size(), [&](std::size_t k) { return qi(S[k]->desired_used_flow_U_tilde().get_quantity()); });
    put_f("sto_Ut_v", S.size(), [&](std::size_t k) { return to_float(S[k]->desired_used_flow_U_tilde().get_value()); });
    put_i("sto_demand_q", S.size(), [&](std::size_t k) { return qi(S[k]->purchasing_manager->demand_D().get_quantity()); });
    put_f("sto_demand_v", S.size(), [&](std::size_t k) { return to_float(S[k]->purchasing_manager->demand_D().get_value()); });
    put_f("sto_expected_costs", S.size(), [&](std::size_t k) { return to_float(S[k]->purchasing_manager->expected_costs()); });
    put_f("sto_optimized_value", S.size(), [&](std::size_t k) { return S[k]->purchasing_manager->optimized_value(); });
    put_f("sto_transport_penalty", S.size(), [&](std::size_t k) { return to_float(S[k]->purchasing_manager->total_transport_penalty()); });
    const auto& A = g_init->agents;
    auto firm = [&](std::size_t k) -> const Firm* { return A[k]->is_firm() ? A[k]->as_firm() : nullptr; };
    put_f("agent_forcing", A.size(), [&](std::size_t k) { return A[k]->forcing(); });
    if (g_cfg->F("model_f64")[MP_UTILITARIAN] != 0.0)
        put_f("consumer_utility", A.size(), [&](std::size_t k) { return A[k]->is_consumer() ? A[k]->as_consumer()->utility : 0.0; });
    put_i("firm_X_q", A.size(), [&](std::size_t k) { return firm(k) ? qi(firm(k)->production_X().get_quantity()) : 0; });
    put_f("firm_X_v", A.size(), [&](std::size_t k) { return firm(k) ? to_float(firm(k)->production_X().get_value()) : 0.0; });
    put_i("firm_Xhat_q", A.size(), [&](std::size_t k) { return firm(k) ? qi(firm(k)->sales_manager->communicated_parameters().possible_production_X_hat.get_quantity()) : 0; });
    put_i("firm_Xhat_p", A.size(), [&](std::size_t k) { return firm(k) ? pi(firm(k)->sales_manager->communicated_parameters().possible_production_X_hat.get_price()) : 0; });
    put_i("firm_Xexp_q", A.size(), [&](std::size_t k) { return firm(k) ? qi(firm(k)->sales_manager->communicated_parameters().expected_production_X.get_quantity()) : 0; });
    put_f("firm_Xexp_v", A.size(), [&](std::size_t k) { return firm(k) ? to_float(firm(k)->sales_manager->communicated_parameters().expected_production_X.get_value()) : 0.0; });
    put_i("firm_nbar", A.size(), [&](std::size_t k) { return firm(k) ? pi(firm(k)->sales_manager->communicated_parameters().offer_price_n_bar) : 0; });
    put_f("firm_costs_C", A.size(), [&](std::size_t k) { return firm(k) ? to_float(firm(k)->sales_manager->total_production_costs_C()) : 0.0; });
    put_f("firm_revenue_R", A.size(), [&](std::size_t k) { return firm(k) ? to_float(firm(k)->sales_manager->total_revenue_R()) : 0.0; });
    put_i("firm_desired_q", A.size(), [&](std::size_t k) { return firm(k) ? qi(firm(k)->capacity_manager->desired_production_X_tilde().get_quantity()) : 0; });
    put_i("firm_possible_q", A.size(), [&](std::size_t k) { return firm(k) ? qi(firm(k)->capacity_manager->possible_production_X_hat().get_quantity()) : 0; });
    const std::size_t ns = model->sectors.size(), nr = model->regions.size();
    put_i("sector_production_q", ns, [&](std::size_t k) { return qi(model->sectors[k]->total_production_X().get_quantity()); });
    put_f("sector_production_v", ns, [&](std::size_t k) { return to_float(model->sectors[k]->total_production_X().get_value()); });
    put_i("sector_demand_q", ns, [&](std::size_t k) { return qi(model->sectors[k]->total_demand_D().get_quantity()); });
    put_f("sector_demand_v", ns, [&](std::size_t k) { return to_float(model->sectors[k]->total_demand_D().get_value()); });
    put_i("region_import_q", nr, [&](std::size_t k) { return qi(model->regions[k]->import_flow_Z().get_quantity()); });
    put_f("region_import_v", nr, [&](std::size_t k) { return to_float(model->regions[k]->import_flow_Z().get_value()); });
    put_i("region_export_q", nr, [&](std::size_t k) { return qi(model->regions[k]->export_flow_Z().get_quantity()); });
    put_f("region_export_v", nr, [&](std::size_t k) { return to_float(model->regions[k]->export_flow_Z().get_value()); });
    put_i("region_consumption_q", nr, [&](std::size_t k) { return qi(model->regions[k]->consumption_C().get_quantity()); });
    put_f("region_consumption_v", nr, [&](std::size_t k) { return to_float(model->regions[k]->consumption_C().get_value()); });
    put_f("region_gdp_v", nr, [&](std::size_t k) { return to_float(model->regions[k]->get_gdp().get_value()); });
    if (g_cfg->dump_observables) dump_observables(out, model, P);
}

// the time loop: same order of calls as src/ModelRun.cpp:153-229
void ModelRun::run() {
    if (has_run) throw log::error(this, "Model has already run");
    has_run = true;
    step(IterationStep::INITIALIZATION);
    model_m->time_m = start_time_m;
    model_m->start();
    time_m = 0;
    step(IterationStep::SCENARIO);
    std::unique_ptr<accio::Writer> out;
    if (!g_cfg->dump_path.empty()) out = std::make_unique<accio::Writer>(g_cfg->dump_path);
    while (!done()) {
        g_step = time_m;
        apply_scenario(model_m.get());
        const auto t0 = std::chrono::high_resolution_clock::now();
        model_m->switch_registers();
        step(IterationStep::CONSUMPTION_AND_PRODUCTION);
        model_m->iterate_consumption_and_production();
        step(IterationStep::EXPECTATION);
        model_m->iterate_expectation();
        step(IterationStep::PURCHASE);
        model_m->iterate_purchase();
        step(IterationStep::INVESTMENT);
        model_m->iterate_investment();
        const auto t1 = std::chrono::high_resolution_clock::now();
        g_phase_seconds += std::chrono::duration<double>(t1 - t0).count();
        duration_m = std::chrono::duration_cast<std::chrono::milliseconds>(t1 - t0).count();
        step(IterationStep::OUTPUT);
        if (out && (time_m % g_cfg->dump_every == 0 || time_m + 1 == unsigned(g_cfg->steps))) {
            dump_step(*out, model_m.get(), time_m, g_cfg->dump_edges);
        }
        step(IterationStep::SCENARIO);
        model_m->tick();
        ++time_m;
    }
    if (out) {
        std::vector<std::int64_t> ev;
        std::vector<double> evv;
        for (const auto& e : g_events) {
            ev.push_back(e.step); ev.push_back(e.type); ev.push_back(e.a); ev.push_back(e.b);
            evv.push_back(e.value);
        }
        out->put("events", ev);
        out->put("event_values", evv);
    }
}

}  // namespace acclimate

int main(int argc, char** argv) {
    if (argc < 3) {
        std::fprintf(stderr, "usage: %s <network.acc> <steps> [dump.acc] [--every k] [--no-edges] [--observe]\n", argv[0]);
        return 1;
    }
    try {
        HarnessConfig cfg;
        cfg.net = accio::read_file(argv[1]);
        cfg.steps = std::atoi(argv[2]);
        for (int i = 3; i < argc; ++i) {
            const std::string a = argv[i];
            if (a == "--every" && i + 1 < argc) cfg.dump_every = std::atoi(argv[++i]);
            else if (a == "--no-edges") cfg.dump_edges = false;
            else if (a == "--observe") cfg.dump_observables = true;
            else if (a == "--ingest" && i + 1 < argc) cfg.ingest_out = argv[++i];
            else cfg.dump_path = a;
        }
        settings::SettingsNode node;
        node.cfg = &cfg;
        const auto t0 = std::chrono::high_resolution_clock::now();
        acclimate::ModelRun run(node);
        if (!cfg.ingest_out.empty()) return 0;  // ingest only: the survivors have been written
        const auto t1 = std::chrono::high_resolution_clock::now();
        run.run();
        const double init_s = std::chrono::duration<double>(t1 - t0).count();
        if (std::getenv("ORACLE_SOLVER_STATS") != nullptr) {
            long long st[5];
            nlopt_provider_stats(st);
            std::fprintf(stderr, "solver %s: calls %lld iters %lld evals %lld qp_iters %lld max_iters %lld\n", nlopt_provider_name(), st[0], st[1], st[2], st[3], st[4]);
        }
        std::printf("{\"edges\": %zu, \"steps\": %d, \"threads\": %u, \"init_seconds\": %.6f, \"phase_seconds\": %.6f, \"events\": %zu}\n",
                    acclimate::g_init->edges.size(), cfg.steps, run.thread_count(), init_s, acclimate::g_phase_seconds, g_events.size());
        return 0;
    } catch (const std::exception& ex) {
        std::fprintf(stderr, "oracle harness: %s\n", ex.what());
        return 255;
    }
}
