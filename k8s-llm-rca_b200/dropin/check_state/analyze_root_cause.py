"""Mirror of the graph half of reference check_state/analyze_root_cause.py (:51-115, :155-221):
the state query per node of a path instance, the missing-state branch and the entity-name lookup.
The semantic analyser is any object with add_message / run_assistant / wait_get_last_k_message
(the reference's OpenAI assistant; here normally the recorded-response stub)."""
import neo4j

from common.openai_generic_assistant import OpenAIGenericAssistant

IMPORTANT_FIELDS = ("status", "spec", "path", "server", "subsets", "roleRef", "subjects", "rules", "webhooks",
                    "secrets", "data", "metadata")


def _state_query(entityKind, entityId, lo, hi):
    return ("MATCH (n1:%s)-[r1:HasState]->(n2:%s)\n"
            "WHERE n1.id = '%s'\n"
            "AND r1.tmin <= '%s' AND r1.tmax > '%s'\n"
            "RETURN n2\nLIMIT 10;" % (entityKind, entityKind.upper(), entityId, lo, hi))


def find_loose_states(entityKind, entityId, tmin, tmax):
    """STATE intervals that overlap [tmin, tmax)   [reference :51-60]"""
    return _state_query(entityKind, entityId, tmax, tmin)


def find_strict_states(entityKind, entityId, timestamp):
    """STATE intervals [tmin, tmax) that contain the timestamp   [reference :70-79]"""
    return _state_query(entityKind, entityId, timestamp, timestamp)


def _entity_name_key(entity):
    if entity["isNative"] == "true":
        return "name2"
    if entity["isAtomic"] == "true":
        return "val"
    by_tag = {"nfs": "path", "hostPath": "path", "container": "containerName", "image": "imageName"}
    tag = entity["tag"]
    if tag not in by_tag:
        # the reference's if/elif ladder leaves `key` unassigned for any other tag (:210-221)
        raise UnboundLocalError("cannot access local variable 'key' where it is not associated with a value")
    return by_tag[tag]


def ad_hoc_find_entity_name(entity_kind, entity_id, query_executor):
    """Display name of an entity, by id   [reference :200-221]"""
    records = query_executor.run_query(
        "match (n1:%s)\nwhere n1.id = '%s'\nreturn n1\nlimit 1" % (entity_kind, entity_id))
    entity = records[0]["n1"]
    return entity[_entity_name_key(entity)]


def check_semantic(state_node, error_message, semanticAnalyzer):
    """One LLM round over the important fields of a STATE node   [reference :223-250]"""
    fields = {k: state_node[k] for k in set(state_node.keys()) & set(IMPORTANT_FIELDS)}
    prompt = ("Recorded state of a %s object, as JSON. Check 'spec' and 'status' (other fields when those are absent) for "
              "anything that explains the error message; do not propose a fix.\nError message:\n%s\nJSON:\n%s\n"
              % (state_node["kind"], error_message, fields))
    semanticAnalyzer.add_message(prompt)
    semanticAnalyzer.run_assistant()
    return semanticAnalyzer.wait_get_last_k_message(1).data[0].content[0].text.value


def check_states_existence_and_semantic(query_executor, cypher_query, semanticAnalyzer, error_message):
    """[reference :155-170]"""
    records = query_executor.run_query(cypher_query)
    if len(records) == 0:
        return ["There is not a STATE node corresponds to the Entity node"]
    clues = []
    for rec in records:
        state = rec["n2"]
        clues.append(state["kind"] + "(" + state["id"] + "): " + check_semantic(state, error_message, semanticAnalyzer))
    return clues


def check_states_of_entity(entity_kind, entity_id, error_message, timestamp, query_executor, semanticAnalyzer):
    """State of one entity at the event's time: a 'missing' clue or one clue per STATE   [:173-197]"""
    records = query_executor.run_query(find_strict_states(entity_kind, entity_id, timestamp))
    clues = []
    if len(records) == 0:
        name = ad_hoc_find_entity_name(entity_kind, entity_id, query_executor)
        missing = ("%s (%s): there is not a STATE (%s) node corresponds to the Entity (%s) node, which is an apparent "
                   "error. we confirm that %s does not exist." % (entity_kind, entity_id, entity_kind.upper(), entity_kind, name))
        clues.append(missing)
        semanticAnalyzer.add_message(missing)
    else:
        for rec in records:
            state = rec["n2"]
            clues.append(state["kind"].upper() + "(" + state["id"] + "): " +
                         check_semantic(state, error_message, semanticAnalyzer))
    for clue in clues:
        print("~" * 100 + "\n" + clue + "\n" + "~" * 100)
    return clues


def check_statepath(query_executor, semanticAnalyzer, statepath):
    """One instance row -> (report, {entity: clues})   [reference :82-150]"""
    for ele in statepath:
        if isinstance(ele, neo4j.graph.Node) and ele["kind"] == "Event":
            timestamp, error_message = ele["timestamp"], ele["message"]
    path_clues, kinds = {}, []
    for ele in statepath:
        if not isinstance(ele, neo4j.graph.Node) or ele["kind2"] == "Event" or ele["kind"] == "Event":
            continue
        if ele["isNative"] == "true":
            entity_kind = ele["kind2"]
        elif ele["isNative"] == "false":
            entity_kind = ele["tag"]
        entity_id = ele["id"]
        print(entity_kind, entity_id)
        kinds.append(entity_kind)
        path_clues["%s(%s)" % (entity_kind, entity_id)] = check_states_of_entity(
            entity_kind, entity_id, error_message, timestamp, query_executor, semanticAnalyzer)
    semanticAnalyzer.add_message(
        "Based on the previous analysis of %s, summarize the root cause of the error message, score every kind's "
        "relevance (0~10/10), and give a resolution with kubectl or bash commands where applicable. Answer as JSON with "
        "the keys summary (list of kind / explanation / relevance_score), conclusion and resolution." % ", ".join(kinds))
    semanticAnalyzer.run_assistant()
    report = semanticAnalyzer.wait_get_last_k_message(1).data[0].content[0].text.value
    return report, path_clues


# ---- LLM plumbing (reference :6-46)
def setup_state_semantic_analyzer():
    analyzer = OpenAIGenericAssistant()
    analyzer.create_assistant("You read the state of k8s objects and relate mistakes in it to an error message.",
                              "k8s-state-semantic-analyzer", "gpt-4")
    analyzer.create_thread()
    print(analyzer.assistant.id)
    print(analyzer.thread.id)
    analyzer.add_message("Every entity should have a STATE node for the time of the event; an entity without one does not "
                         "exist or was never created, which is an error in itself.")
    return analyzer
