// oracle/matchshim: empty stand-in for <boost/serialization/access.hpp> (test infrastructure); names come from facade.h
