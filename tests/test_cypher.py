"""Cypher-subset front end: the reference's statements compile to the right plans; bad text raises
the exception the reference's retry loop catches (test_all.py:109)."""
import pytest

from k8s_llm_rca_b200 import cypher as cy

S1 = """
        MATCH (n1)
        WHERE n1.category IN ['NativeEntity', 'ExternalEntity']
        RETURN n1.category AS category, n1.kind AS kind
        """
S2 = """
        MATCH (n1:Event)-[s1:HasEvent]->(N1:EVENT)
        WHERE N1.message contains $message
        WITH n1, N1, s1
        MATCH (n1:Event)-[r1:ReferInternal]->(n2)
        WHERE r1.key = 'involvedObject_uid'
        RETURN distinct n2.kind2
        LIMIT 5;
        """
S3 = """
        MATCH path = (n1)-[*1..3]{arrow}(n2)
        WHERE n1.kind = $srcKind and n2.kind = $destKind
        AND all(node in nodes(path) WHERE single(x in nodes(path) WHERE x = node))
        AND all(node in nodes(path) WHERE not node.kind in ['Event', 'Namespace'])
        AND ($intermediateKinds IS NULL
            OR size($intermediateKinds) = 0
            OR any(node in nodes(path)[1..-1] WHERE node.kind in $intermediateKinds))
        RETURN path
        """
S3C = "MATCH path = (n1)-[r1]-(n2) WHERE n1.kind = $srcKind and n2.kind = $destKind RETURN path"
S3D = ("MATCH path = (n1)-[r1]-(n2)-[r2]-(n3) WHERE n1.kind = $srcKind and n2.kind = 'Namespace' "
       "and n3.kind = $destKind RETURN path")
S5 = """
    MATCH (n1:ResourceQuota)-[r1:HasState]->(n2:RESOURCEQUOTA)
    WHERE n1.id = '66637a34-b552-448f-9da6-976aa7462533'
    AND r1.tmin <= '2020-12-11 06:35:02.011' AND r1.tmax > '2020-12-11 06:35:02.011'
    RETURN n2
    LIMIT 10;
    """
S6 = "match (n1:Pod)\nwhere n1.id = 'abc'\nreturn n1\nlimit 1"
# the worked example inside the reference's LLM template (generate_query.py:197-207)
TEMPLATE_EXAMPLE = """MATCH (evt:EVENT)
    WHERE evt.message CONTAINS 'MountVolume.SetUp failed for volume "gen-white-list-conf" : configmap "es-gen-white-list-configmap" not found'
    WITH evt
    LIMIT 1
    MATCH (event:Event)-[r1:HasEvent]->(evt)
    WHERE r1.key = 'metadata_uid'
    MATCH (event)-[r2:ReferInternal]->(pod:Pod)
    WHERE r2.key = 'involvedObject_uid'
    MATCH (pod)-[r3:ReferInternal]->(configMap:ConfigMap)
    WHERE r3.key = 'spec_volumes_configMap_name'
    RETURN event, r1, evt, r2, pod, r3, configMap"""
P3 = {"srcKind": "Pod", "destKind": "nfs", "intermediateKinds": ["PersistentVolume"]}


def test_kind_listing_is_a_node_scan():
    p = cy.compile_query(S1)
    assert isinstance(p, cy.NodeScanPlan) and p.label is None and p.id_equals is None
    assert [a for (_, a, _) in p.ret.items] == ["category", "kind"]


def test_src_kind_query_is_a_chain_from_the_event_state():
    p = cy.compile_query(S2, {"message": "x"})
    assert isinstance(p, cy.ChainPlan)
    assert p.seed_var == "N1" and p.walk == [("N1", "n1", "in"), ("n1", "n2", "out")]
    assert [s.key for s in p.steps] == [None, "involvedObject_uid"]
    assert p.labels == {"n1": "Event", "N1": "EVENT", "n2": None}
    assert p.ret.distinct and p.ret.limit == 5 and p.ret.items[0][2] == "n2.kind2"     # column named by its text


@pytest.mark.parametrize("arrow,direction", [("->", "out"), ("-", "und")])
def test_metapath_queries(arrow, direction):
    p = cy.compile_query(S3.format(arrow=arrow), P3)
    assert isinstance(p, cy.VarLenPlan)
    assert (p.direction, p.min_hops, p.max_hops) == (direction, 1, 3)
    assert (p.src_kind, p.dest_kind, p.kind_prop) == ("Pod", "nfs", "kind")
    assert p.node_simple and p.excl_kinds == ["Event", "Namespace"] and p.inter_kinds == ["PersistentVolume"]
    for empty in ([], None):
        q = cy.compile_query(S3.format(arrow=arrow), dict(P3, intermediateKinds=empty))
        assert q.inter_kinds == []            # null or empty list: the any(...) rule is vacuous


def test_hop_bound_comes_from_the_text():
    p = cy.compile_query(S3.format(arrow="-").replace("*1..3", "*1..5"), P3)
    assert p.max_hops == 5


def test_single_hop_and_namespace_rungs():
    c = cy.compile_query(S3C, P3)
    assert (c.min_hops, c.max_hops, c.direction, c.pos_kinds, c.node_simple) == (1, 1, "und", ["Pod", "nfs"], False)
    d = cy.compile_query(S3D, P3)
    assert (d.min_hops, d.max_hops, d.pos_kinds, d.node_simple) == (2, 2, ["Pod", "Namespace", "nfs"], False)
    assert d.rel_vars == ["r1", "r2"]


def test_state_queries():
    p = cy.compile_query(S5)
    assert isinstance(p, cy.StatePlan)
    assert (p.entity_label, p.state_label, p.entity_id) == ("ResourceQuota", "RESOURCEQUOTA", "66637a34-b552-448f-9da6-976aa7462533")
    assert p.lo == p.hi == "2020-12-11 06:35:02.011" and p.ret.limit == 10
    loose = S5.replace("r1.tmin <= '2020-12-11 06:35:02.011'", "r1.tmin <= '2020-12-11 07:00:00.000'")
    q = cy.compile_query(loose)
    assert (q.lo, q.hi) == ("2020-12-11 07:00:00.000", "2020-12-11 06:35:02.011")


def test_entity_lookup_lower_case_keywords():
    p = cy.compile_query(S6)
    assert isinstance(p, cy.NodeScanPlan) and p.label == "Pod" and p.id_equals == "abc" and p.ret.limit == 1


def test_llm_template_example_is_the_same_chain_as_the_generated_query():
    p = cy.compile_query(TEMPLATE_EXAMPLE)
    assert isinstance(p, cy.ChainPlan) and p.seed_var == "evt" and p.seed_limit == 1
    assert p.walk == [("evt", "event", "in"), ("event", "pod", "out"), ("pod", "configMap", "out")]
    assert p.labels == {"evt": "EVENT", "event": "Event", "pod": "Pod", "configMap": "ConfigMap"}
    assert [s.key for s in p.steps] == ["metadata_uid", "involvedObject_uid", "spec_volumes_configMap_name"]
    assert [src for (_, _, src) in p.ret.items] == ["event", "r1", "evt", "r2", "pod", "r3", "configMap"]


def test_inline_property_maps_and_left_arrows():
    q = ("MATCH (evt:EVENT) WHERE evt.message CONTAINS 'x' WITH evt LIMIT 1 "
         "MATCH (evt)<-[r1:HasEvent {key: 'metadata_uid'}]-(e:Event) "
         "MATCH (e)-[r2:ReferInternal {key: 'involvedObject_uid'}]->(p:Pod) RETURN p, r2, e, r1, evt")
    p = cy.compile_query(q)
    assert p.walk == [("evt", "e", "in"), ("e", "p", "out")] and [s.key for s in p.steps] == ["metadata_uid", "involvedObject_uid"]


def test_string_literals_round_trip_python_repr():
    for msg in ['plain', 'with "double" quotes', "with 'single' quotes", "both ' and \"", "back\\slash and \n newline"]:
        q = f"MATCH (evt:EVENT) WHERE evt.message CONTAINS {msg!r} WITH evt LIMIT 1 MATCH (e:Event)-[r:HasEvent]->(evt) RETURN e"
        p = cy.compile_query(q)
        assert p.seed_where[3] == ("lit", msg)


@pytest.mark.parametrize("bad", [
    "", "MATCH (n RETURN n", "MATC (n) RETURN n", "MATCH (n)-[r]->(m) WHERE n.x = 1", "MATCH (n) RETURN",
    "MATCH (n) WHERE n.a = RETURN n", "CREATE (n) RETURN n", "MATCH (a)-[r:T]->(b) DELETE r",
    "MATCH (a)<-[r:T]->(b) RETURN a", "MATCH (n) WHERE n.x = 'unterminated RETURN n",
    "MATCH path = (a)-[*]->(b) WHERE a.kind = 'x' RETURN path",
    "MATCH (a:K)-[r:T]->(b:K), (c) RETURN a",
])
def test_bad_or_unsupported_text_raises_cypher_syntax_error(bad):
    with pytest.raises(cy.CypherSyntaxError):
        cy.compile_query(bad)


def test_missing_parameter_is_reported():
    with pytest.raises(cy.CypherSyntaxError) as ei:
        cy.compile_query(S3.format(arrow="-"), {"srcKind": "Pod"})
    assert "destKind" in str(ei.value) or "parameter" in str(ei.value).lower()


def test_three_valued_logic():
    env = {"n": {"a": 1}}
    ev = lambda t: cy.eval_expr(cy.Parser(t).expr(), env, {})
    assert ev("n.a = 1") is True and ev("n.missing = 1") is None and ev("not n.missing = 1") is None
    assert ev("n.missing = 1 or n.a = 1") is True and ev("n.missing = 1 and n.a = 2") is False
    assert ev("n.missing in ['x']") is None and ev("n.missing is null") is True


def test_shared_parse_lifts_string_literals_and_falls_back_when_unsafe():
    """parse_shared: statements that differ only in their string literals share one parse; the literals come
    back as parameters; anything unsafe (a literal in RETURN, bad text) is parsed as written."""
    a = S5
    b = S5.replace("66637a34-b552-448f-9da6-976aa7462533", "0000").replace("06:35:02.011", "07:00:00.000")
    ca, la = cy.parse_shared(a)
    cb, lb = cy.parse_shared(b)
    assert ca is cb                                            # one skeleton, parsed once
    assert la == {"__s0": "66637a34-b552-448f-9da6-976aa7462533", "__s1": "2020-12-11 06:35:02.011",
                  "__s2": "2020-12-11 06:35:02.011"}
    assert lb["__s0"] == "0000" and lb["__s1"] == "2020-12-11 07:00:00.000"
    # the compiled plan asks for the same things either way
    bind = {}
    lifted, plain = cy.compile_query(a, None, bind=bind), cy.compile_query(a)
    assert type(lifted) is type(plain) and bind == la
    assert cy.eval_expr(("param", "__s0"), {}, bind) == "66637a34-b552-448f-9da6-976aa7462533"
    # quotes and escapes inside literals survive the lifting
    for msg in ['with "double" quotes', "with 'single' quotes", "both ' and \"", "back\\slash and \n newline", "$__not a param"]:
        q = f"MATCH (evt:EVENT) WHERE evt.message CONTAINS {msg!r} WITH evt LIMIT 1 MATCH (e:Event)-[r:HasEvent]->(evt) RETURN e"
        bind = {}
        p = cy.compile_query(q, None, bind=bind)
        assert cy.eval_expr(p.seed_where[3], {}, bind) == msg
    # a literal in RETURN names its column: not lifted
    c, l = cy.parse_shared("MATCH (n:Pod) WHERE n.id = 'x' RETURN 'tag', n")
    assert l == {} and c[-1].items[0][2] == "'tag'"
    # errors quote the statement as written, not the skeleton
    with pytest.raises(cy.CypherSyntaxError) as ei:
        cy.compile_query("MATCH (n:Pod) WHERE n.id = 'abc' RETURN", None, bind={})
    assert "__s0" not in str(ei.value)


def test_plans_are_kept_per_skeleton_and_per_value_the_compilation_looked_at():
    """Statements that differ in the message alone share one compiled plan; a different relationship key is a
    different plan; a state statement is bound to its own id and timestamps every time."""
    def s4(msg, key):
        return TEMPLATE_EXAMPLE.replace("spec_volumes_configMap_name", key).replace('configmap "es-gen-white-list-configmap" not found', msg)
    b1, b2, b3 = {}, {}, {}
    p1 = cy.compile_query(s4("first message", "spec_volumes_configMap_name"), {}, bind=b1)
    p2 = cy.compile_query(s4("second message", "spec_volumes_configMap_name"), {}, bind=b2)
    p3 = cy.compile_query(s4("second message", "spec_containers_envFrom_configMapRef_name"), {}, bind=b3)
    assert p1 is p2 and p3 is not p1
    assert [s.key for s in p3.steps][-1] == "spec_containers_envFrom_configMapRef_name"
    assert [s.key for s in p1.steps][-1] == "spec_volumes_configMap_name"
    # the message stays an expression of the shared plan, evaluated with each statement's own literals
    assert cy.eval_expr(p1.seed_where[3], {}, b1) != cy.eval_expr(p2.seed_where[3], {}, b2)
    assert "second message" in cy.eval_expr(p2.seed_where[3], {}, b2)
    q = S5.replace("66637a34-b552-448f-9da6-976aa7462533", "{id}").replace("2020-12-11 06:35:02.011", "{ts}")
    a = cy.compile_query(q.format(id="a" * 8, ts="2020-12-11 06:35:02.011"), {}, bind={})
    b = cy.compile_query(q.format(id="b" * 8, ts="2021-01-01 00:00:00.000"), {}, bind={})
    assert (a.entity_id, a.lo, a.hi) == ("a" * 8, "2020-12-11 06:35:02.011", "2020-12-11 06:35:02.011")
    assert (b.entity_id, b.lo, b.hi) == ("b" * 8, "2021-01-01 00:00:00.000", "2021-01-01 00:00:00.000")
    # parameters the caller passes count like literals: another kind, another plan
    v1 = cy.compile_query(S3.format(arrow="-"), {"srcKind": "Pod", "destKind": "nfs", "intermediateKinds": []}, bind={})
    v2 = cy.compile_query(S3.format(arrow="-"), {"srcKind": "Pod", "destKind": "Node", "intermediateKinds": []}, bind={})
    v3 = cy.compile_query(S3.format(arrow="-"), {"srcKind": "Pod", "destKind": "Node", "intermediateKinds": ["PersistentVolume"]}, bind={})
    assert (v1.dest_kind, v2.dest_kind, v3.inter_kinds) == ("nfs", "Node", ["PersistentVolume"]) and v2.inter_kinds == []
    # a missing parameter is still reported, however often the skeleton was compiled before
    with pytest.raises(cy.CypherSyntaxError):
        cy.compile_query(S3.format(arrow="-"), {"srcKind": "Pod"}, bind={})


def test_a_quote_inside_a_comment_is_not_a_string_literal():
    q = "/* the pod's claim */ MATCH (n1:Pod) WHERE n1.id = 'abc' // it's here\nRETURN n1 LIMIT 1"
    p = cy.compile_query(q, {}, bind={})
    assert isinstance(p, cy.NodeScanPlan) and p.id_equals == "abc"


def test_a_url_in_a_literal_still_shares_its_plan():
    q = "MATCH (evt:EVENT) WHERE evt.message CONTAINS 'pull http://registry/{}' WITH evt LIMIT 1 MATCH (e:Event)-[r:HasEvent]->(evt) RETURN e"
    a, b = cy.compile_query(q.format("a"), {}, bind={}), cy.compile_query(q.format("b"), {}, bind={})
    assert a is b
